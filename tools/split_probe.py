"""Tuning aid: sharded.prune_tensor_split on one big conv tensor under torchrun (or alone), next to the single-GPU plan.
python -m torch.distributed.run --nproc-per-node N tools/split_probe.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from wavelettransforms_b200 import _lib, engine, sharded

rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if world > 1:
    dist.init_process_group('nccl', rank=rank, world_size=world)
g = torch.Generator(device=dev).manual_seed(1000)
for shape, wavelet, level in (((2048, 512, 3, 3), 'bior4.4', 5), ((2048, 512, 3, 3), 'haar', 1), ((4096, 4096, 1, 1), 'bior2.2', 3)):
    w = torch.randn(shape, device=dev, generator=g) * 0.02
    for gather in (False, True):
        for _ in range(2):
            sharded.prune_tensor_split(w, wavelet, level, 90.0, gather=gather)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(5):
            out, t, zeros = sharded.prune_tensor_split(w, wavelet, level, 90.0, gather=gather)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = (time.perf_counter() - t0) / 5 * 1e3
        if rank == 0:
            print(f'{shape} {wavelet} L{level} world {world} gather={gather}: {ms:.3f} ms/call (wall, plan creation included)', flush=True)
    if rank == 0:
        for _ in range(2):
            engine.prune_device_tensors([w], [level], wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, 90.0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            engine.prune_device_tensors([w], [level], wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, 90.0)
        torch.cuda.synchronize()
        print(f'{shape} {wavelet} L{level} single-GPU one-shot plan: {(time.perf_counter() - t0) / 5 * 1e3:.3f} ms/call', flush=True)
    if world > 1:
        dist.barrier()
if world > 1:
    dist.destroy_process_group()
