"""Probe: does torch's symmetric memory (CUDA VMM peer mappings) come up between the GPUs of this box?
torchrun --nproc-per-node 2 tools/symm_probe.py"""
import os
import time

import torch
import torch.distributed as dist

rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
dev = torch.device('cuda', int(os.environ['LOCAL_RANK']))
dist.init_process_group('nccl', device_id=dev)
import torch.distributed._symmetric_memory as symm_mem

n = 64 << 20
t = symm_mem.empty(n, dtype=torch.float32, device=dev)
hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
print(rank, 'ptrs', [hex(p) for p in hdl.buffer_ptrs], flush=True)
t.fill_(float(rank))
hdl.barrier()
peer = hdl.get_buffer((rank + 1) % world, (n,), torch.float32)
src = torch.full((n // world,), 100.0 + rank, device=dev)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(3):
    peer[rank * (n // world):(rank + 1) * (n // world)].copy_(src)
e0.record()
for _ in range(10):
    peer[rank * (n // world):(rank + 1) * (n // world)].copy_(src)
e1.record()
torch.cuda.synchronize()
hdl.barrier()
ms = e0.elapsed_time(e1) / 10
other = (rank - 1) % world
got = t[other * (n // world)].item()
print(rank, 'peer store of', src.numel() * 4 / 1e6, 'MB:', ms, 'ms ->', src.numel() * 4 / ms / 1e6, 'GB/s; saw', got, 'want', 100.0 + other, flush=True)
dist.destroy_process_group()
