"""Layer-sharded DWT pruning of one model over the GPUs of a box.

Every tensor's transform, per-tensor percentile and reconstruction are independent of every other
tensor (/root/reference/ResNet/dwt_pruning.py:155-160 prunes layer by layer), so the conv layers are
dealt to ranks by a size-balanced greedy (workloads.lpt_assignment), each rank prunes its own layers
in ONE grouped plan whose outputs land directly in its slot of a flat, padded send buffer (the pack
step is the kernels' own store), and ONE all-gather over NCCL / NVLink reassembles the pruned
state on every rank; a second, tiny all-gather carries the int64 per-layer counts.

The process group is the caller's (`torch.distributed`, backend nccl on the GPU box).  `prune_fn`
is injectable so the host logic (assignment, packing, reassembly) is testable on CPU with gloo and
the oracle standing in for the kernels (tests/test_sharded.py); the default runs libwtprune.
"""
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

from . import workloads

PruneFn = Callable[[List[torch.Tensor], List[torch.Tensor], List[int]], np.ndarray]


def _gpu_prune_fn(wavelet, mode, threshold_kind, threshold_value, numpy_semantics) -> PruneFn:
    from . import engine

    def fn(xs, outs, levels):
        res = engine.prune_device_tensors(xs, levels, wavelet, mode, threshold_kind, threshold_value, numpy_semantics,
                                          outs=outs)
        c = np.zeros((len(xs), 4), dtype=np.int64)
        c[:, :3] = res.counts
        c[:, 3] = res.nonzero_before
        return c
    return fn


class ShardLayout:
    """Who owns which layer and where every layer lives in the gathered flat buffer."""

    def __init__(self, shapes: Sequence[Tuple[int, ...]], world_size: int, align: int = 64):
        self.shapes = [tuple(int(d) for d in s) for s in shapes]
        self.numels = [int(np.prod(s, dtype=np.int64)) for s in self.shapes]
        self.world_size = world_size
        self.owner = workloads.lpt_assignment(self.numels, world_size)
        self.offset = [0] * len(self.shapes)      # element offset inside the owner's shard
        fill = [0] * world_size
        for i, (r, n) in enumerate(zip(self.owner, self.numels)):
            self.offset[i] = fill[r]
            fill[r] += (n + align - 1) // align * align     # 256-byte aligned slots: vector loads / TMA stay legal
        self.shard_elems = max(max(fill), align)
        self.load = [sum(n for n, o in zip(self.numels, self.owner) if o == r) for r in range(world_size)]

    def owned(self, rank: int) -> List[int]:
        return [i for i, o in enumerate(self.owner) if o == rank]

    @property
    def balance(self) -> float:
        """ideal / max shard load: the ceiling on strong-scaling efficiency (ResNet-18 at 8 ranks: 0.59)"""
        return (sum(self.numels) / self.world_size) / max(max(self.load), 1)

    def view(self, flat: torch.Tensor, i: int) -> torch.Tensor:
        base = self.owner[i] * self.shard_elems + self.offset[i]
        return flat[base:base + self.numels[i]].view(self.shapes[i])


def prune_sharded(weights: List[torch.Tensor], levels: Sequence[int], prune_fn: PruneFn,
                  group: Optional[dist.ProcessGroup] = None, layout: Optional[ShardLayout] = None,
                  gather: bool = True):
    """weights: the full (replicated) list on this rank's device.  Returns (pruned tensors -- views into
    the gathered flat buffer, every layer present on every rank --, counts int64 [n_layers, 4]:
    numel, nonzero after, zeros after, nonzero before; flat buffer; layout).  gather=False skips both
    collectives (prune-only timing): then only the owned layers' views are meaningful."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    layout = layout or ShardLayout([w.shape for w in weights], world)
    dev = weights[0].device if weights else torch.device('cpu')
    flat = torch.zeros(world * layout.shard_elems, dtype=torch.float32, device=dev)
    mine = layout.owned(rank)
    my_shard = flat[rank * layout.shard_elems:(rank + 1) * layout.shard_elems]
    xs = [weights[i].contiguous() for i in mine]
    outs = [layout.view(flat, i) for i in mine]          # kernels store straight into the send slot
    counts_local = torch.zeros((len(weights), 4), dtype=torch.int64)
    if mine:
        c = prune_fn(xs, outs, [int(levels[i]) for i in mine])
        counts_local[mine] = torch.from_numpy(np.asarray(c, dtype=np.int64))
    counts = counts_local.to(dev)
    if gather and world > 1:
        dist.all_gather_into_tensor(flat, my_shard.clone() if dev.type == 'cpu' else my_shard, group=group)
        # every layer has exactly one owner and the others contribute zeros: a sum is the gather of the counts
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    pruned = [layout.view(flat, i) for i in range(len(weights))]
    return pruned, counts.cpu().numpy(), flat, layout


def prune_model_sharded(weights: List[torch.Tensor], wavelet: str, level: int, percentile: float, *,
                        mode: str = 'periodization', clamp_level: bool = True, numpy_semantics: Optional[str] = None,
                        group: Optional[dist.ProcessGroup] = None, gather: bool = True):
    """dwt_pruning.wavelet_pruning's per-layer loop (dwt_pruning.py:155-164), layer-sharded.  Each layer is
    its own multi_resolution_analysis call in the reference, so the clamp (:64-65) is per layer."""
    from . import _lib, config
    _, F = _lib.wavelet_lookup(wavelet)
    if level < 0:
        raise ValueError(f"Level value of {level} is too low . Minimum level is 0.")
    levels = []
    for w in weights:
        if w.ndim < 2:
            levels.append(0)
        else:
            levels.append(min(level, _lib.dwt_max_level(min(w.shape[-2:]), F)) if clamp_level else level)
    fn = _gpu_prune_fn(wavelet, mode, _lib.THRESHOLD_PERCENTILE, float(percentile), numpy_semantics or config.numpy_semantics)
    return prune_sharded(weights, levels, fn, group=group, gather=gather)
