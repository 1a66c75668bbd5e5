"""Tuning aid: e2e (host buffers in/out through the host-I/O plan) for several pipeline depths, next to the raw
pinned-memory copy times of the same bytes."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

name = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
dev = torch.device('cuda', 0)
torch.cuda.set_device(0)
from wavelettransforms_b200 import workloads
census = workloads.conv_census(bench.WORKLOADS[name][0])
host = workloads.random_init_weights(census, seed=0, pin_memory=True)
devt = [torch.empty_like(h, device=dev) for h in host]
back = [torch.empty_like(h).pin_memory() for h in host]
s2 = torch.cuda.Stream()
for mode in ('h2d', 'd2h', 'both'):
    for _ in range(3):
        t0 = time.perf_counter()
        if mode in ('h2d', 'both'):
            for h, d in zip(host, devt):
                d.copy_(h, non_blocking=True)
        if mode in ('d2h', 'both'):
            with torch.cuda.stream(s2):
                for b, d in zip(back, devt):
                    b.copy_(d, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    print(mode, f'{dt * 1e3:.3f} ms')
for g in (1, 2, 3, 4, 6, 8, 12):
    os.environ['WT_PIPE_GROUPS'] = str(g)
    ms, *_ = bench.time_e2e(name, 30, 5, 1, dev)
    print('groups', g, f'{ms / 30:.3f} ms/step')
