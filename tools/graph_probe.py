"""Device time per step of one plan per workload, plain launches against CUDA-graph replay: python tools/graph_probe.py [workload ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench

if __name__ == '__main__':
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(0)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for name in sys.argv[1:] or [bench.DEFAULT_WORKLOAD]:
            out = []
            for graph in (False, True, False, True):
                wl = bench.DeviceWorkload(name, 4, dev)
                for p in wl.plans:
                    p.set_graph(graph)
                ms, _ = bench.time_device_steps(wl, 200, 20, 1, dev)
                out.append(f"{'graph' if graph else 'plain'}={ms / 200 * 1e3:.1f}")
                wl.close()
            print(name, ' '.join(out), 'us/step', flush=True)
