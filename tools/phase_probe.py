"""Per-launch-group device times of one plan per workload (wt_plan_set_timing): python tools/phase_probe.py [workload ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench

if __name__ == '__main__':
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(0)
    for name in sys.argv[1:] or [bench.DEFAULT_WORKLOAD]:
        wl = bench.DeviceWorkload(name, 4, dev)
        bench.time_device_steps(wl, 10, 3, 1, dev)
        ph = bench.phase_profile(wl, 20)
        print(name, wl.plans[0].select_path, ' '.join(f'{n}={t * 1e3:.1f}' for n, t, _ in ph), f'sum={sum(t for _, t, _ in ph) * 1e3:.1f} us', flush=True)
        wl.close()
