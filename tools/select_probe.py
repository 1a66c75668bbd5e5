"""Tuning aid: runs one workload of bench.py a few times with WT_SELECT_STAMPS=1 so libwtprune prints the
phase boundaries of the select kernel (CTA 0's view), plus the per-launch-group cudaEvent times."""
import os
import sys

os.environ['WT_SELECT_STAMPS'] = '1'
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

name = sys.argv[1] if len(sys.argv) > 1 else 'resnet18_bior4.4_L5_p50_clamped'
dev = torch.device('cuda', 0)
wl = bench.DeviceWorkload(name, 4, dev)
stream = int(torch.cuda.current_stream().cuda_stream)
for i in range(6):
    wl.step(i, stream)
    wl.plans[i % 4].results(stream)
print(bench.phase_profile(wl, 20))
