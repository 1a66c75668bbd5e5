"""CPU baselines of BASELINE.md section 4, measured on the box this runs on: the oracle's restatement of the
reference flow (oracle/reference_flow.py: C DWT + live numpy percentile/where + live torch topk/randperm) on ONE
core and on all host cores (process pool over layers).  Bounded samples; prints one JSON line per config."""
import json
import multiprocessing as mp
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import oracle
from wavelettransforms_b200 import workloads

BUDGET_S = float(os.environ.get('CPU_BASELINE_SECONDS', '10'))


def _dwt_job(args):
    shape, arr, wavelet, level, pct, clamp = args
    torch.set_num_threads(1)
    from oracle import reference_flow as rf
    w = torch.from_numpy(arr).view(shape)
    t0 = time.perf_counter()
    rf.multi_resolution_analysis([w], wavelet, level, pct, numpy_semantics='live', clamp_level=clamp)
    return time.perf_counter() - t0


def _mask_job(args):
    shape, arr, kind = args
    torch.set_num_threads(1)
    from oracle import reference_flow as rf
    w = torch.from_numpy(arr.copy()).view(shape)
    t0 = time.perf_counter()
    if kind == 'min_weight':
        rf.percentage_min_pruning(w, 0.5)
    else:
        rf.random_prune_(w, w.numel() // 2)
    return time.perf_counter() - t0


def measure(name, tensors, fn, make_args):
    oracle.build()
    torch.set_num_threads(1)
    jobs = [make_args(t) for t in tensors]
    fn(jobs[0])
    spent, elems, n = 0.0, 0, 0
    for t, j in sorted(zip(tensors, jobs), key=lambda tj: tj[0].numel()):
        if spent >= BUDGET_S:
            break
        spent += fn(j)
        elems += t.numel()
        n += 1
    one = elems / spent
    cores = os.cpu_count() or 1
    nproc = min(cores, len(jobs))
    total = sum(t.numel() for t in tensors)
    reps = max(1, int(BUDGET_S * one * nproc / max(total, 1)))
    reps = min(reps, 20)
    with mp.get_context('fork').Pool(nproc) as pool:
        pool.map(fn, jobs[:nproc], chunksize=1)
        t0 = time.perf_counter()
        for _ in range(reps):
            pool.map(fn, sorted(jobs, key=lambda j: -int(np.prod(j[0]))), chunksize=1)
        dt = time.perf_counter() - t0
    print(json.dumps({'config': name, 'elems_per_s_1core': one, 'sample_1core': f'{n} of {len(tensors)} tensors, {spent:.1f} s',
                      'elems_per_s_all_cores': total * reps / dt, 'cores': cores, 'pool': nproc, 'reps': reps,
                      'library': f'oracle C DWT + numpy {np.__version__} + torch {torch.__version__} (PyWavelets absent)'}), flush=True)


def main():
    r18 = workloads.random_init_weights(workloads.conv_census('resnet18'), seed=0)
    dwt = lambda wavelet, level, clamp: (lambda t: (tuple(t.shape), t.numpy(), wavelet, level, 50.0, clamp))
    measure('resnet18 dwt_pruning bior1.3 L1 p50 (clamps to level 0)', r18, _dwt_job, dwt('bior1.3', 1, True))
    measure('resnet18 dwt_pruning haar L5 p50 (clamped: 3x3 -> L1, 7x7 -> L2)', r18, _dwt_job, dwt('haar', 5, True))
    measure('resnet18 bior2.2 L3 p50 unclamped', r18, _dwt_job, dwt('bior2.2', 3, False))
    measure('resnet18 bior4.4 L5 p50 unclamped (bench headline)', r18, _dwt_job, dwt('bior4.4', 5, False))
    g = torch.Generator().manual_seed(1000)
    big = [torch.randn(2048, 2048, generator=g) * 0.02 for _ in range(2)]
    measure('2 x 2048x2048 db4 L5 p50 (quarter of a 4096x4096 plane each)', big, _dwt_job, dwt('db4', 5, True))
    mask = lambda kind: (lambda t: (tuple(t.shape), t.numpy(), kind))
    measure('resnet18 min_weight 50 %', r18, _mask_job, mask('min_weight'))
    measure('resnet18 random 50 %', r18, _mask_job, mask('random'))


if __name__ == '__main__':
    main()
