"""Tuning aid: device time of a few plans under different WT_SEL_TAIL_MAX / WT_SELECT_PATH settings (both are read
at every launch of the select kernel)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from wavelettransforms_b200 import _lib, engine, workloads

if os.environ.get('WT_PROBE_LIB'):  # A/B of two builds of the library
    _lib.LIB_PATH = os.environ['WT_PROBE_LIB']
dev = torch.device('cuda', 0)
stream = int(torch.cuda.current_stream().cuda_stream)


def model_plan(name):
    arch, wavelet, level, pct, clamp = bench.WORKLOADS[name]
    census = workloads.conv_census(arch)
    levels = bench.effective_levels(census, wavelet, level, clamp)
    host = workloads.random_init_weights(census, seed=0)
    xs = [h.to(dev) for h in host]
    outs = [torch.empty_like(x) for x in xs]
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,) for x, o, lv in zip(xs, outs, levels)]
    return _lib.Plan(items, wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, pct, bench.NUMPY_SEMANTICS), (xs, outs)


def bank_plan():
    g = torch.Generator(device=dev).manual_seed(1000)
    xs = [torch.randn((512, 512, 3, 3), device=dev, generator=g) * 0.02 for _ in range(32)]
    outs = [torch.empty_like(x) for x in xs]
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (0,) for x, o in zip(xs, outs)]
    return _lib.Plan(items, 'db4', 'periodization', _lib.THRESHOLD_PERCENTILE, 50.0, bench.NUMPY_SEMANTICS), (xs, outs)


def timeit(plan, n=20):
    for _ in range(3):
        plan.run(stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        plan.run(stream)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


names = sys.argv[1:] or ['bank', 'resnet152_bior2.2_L3_p50_unclamped', 'resnet18_bior4.4_L5_p50_unclamped',
                         'resnet18_bior4.4_L5_p50_clamped', 'resnet18_bior2.2_L3_p50_unclamped', 'resnet18_haar_L5_p50_clamped']
for name in names:
    plan, keep = bank_plan() if name == 'bank' else model_plan(name)
    row = []
    for path in ('', 'radix', 'sampled'):
        for tail in os.environ.get('WT_PROBE_TAILS', '0,16384,32768,65536').split(','):
            if path == 'radix' and tail != '0':
                continue
            os.environ.pop('WT_SELECT_PATH', None)
            if path:
                os.environ['WT_SELECT_PATH'] = path
            os.environ['WT_SEL_TAIL_MAX'] = tail
            row.append(f'{path or "auto"}/{tail}: {timeit(plan) * 1e3:.1f}')
    print(name, ' | '.join(row), flush=True)
    plan.close()
    del keep
    torch.cuda.empty_cache()
