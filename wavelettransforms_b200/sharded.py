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

PruneFn = Callable[..., np.ndarray]     # (xs, outs, levels[, peer_deltas]) -> counts [n, 4]


def _gpu_prune_fn(wavelet, mode, threshold_kind, threshold_value, numpy_semantics) -> PruneFn:
    from . import engine

    def fn(xs, outs, levels, peer_deltas=None):
        res = engine.prune_device_tensors(xs, levels, wavelet, mode, threshold_kind, threshold_value, numpy_semantics,
                                          outs=outs, peer_deltas=peer_deltas)
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


class PeerBuffer:
    """The flat [world][shard_elems] fp32 buffer of a ShardLayout, allocated in symmetric memory (CUDA VMM, every
    rank's copy mapped into every peer over NVLink) so the pruning kernels can store each pruned weight into all
    ranks' buffers themselves (`Plan.set_peers`): the all-gather happens under the arithmetic, tile by tile, and
    what is left of the collective is one barrier."""

    def __init__(self, layout: ShardLayout, device, group: Optional[dist.ProcessGroup] = None):
        import torch.distributed._symmetric_memory as symm_mem
        group = group or dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.flat = symm_mem.empty(self.world * layout.shard_elems, dtype=torch.float32, device=device)
        self.handle = symm_mem.rendezvous(self.flat, group.group_name)
        mine = int(self.handle.buffer_ptrs[self.rank])
        if mine != self.flat.data_ptr():
            raise RuntimeError('symmetric memory: the local mapping is not the tensor itself')
        # byte distance from this rank's buffer to each peer's mapping of its own
        self.deltas = [int(p) - mine for r, p in enumerate(self.handle.buffer_ptrs) if r != self.rank]
        self.flat.zero_()               # slot padding reads as zero, like the torch.zeros buffer of the NCCL mode
        torch.cuda.current_stream().synchronize()
        self.handle.barrier()           # nobody stores into a peer that is still clearing

    def barrier(self):
        """device-side barrier on the current stream (signal pads): every rank's stores have landed after it"""
        self.handle.barrier()


def prune_sharded(weights: List[torch.Tensor], levels: Sequence[int], prune_fn: PruneFn,
                  group: Optional[dist.ProcessGroup] = None, layout: Optional[ShardLayout] = None,
                  gather=True, peer_buffer: Optional[PeerBuffer] = None):
    """weights: the full (replicated) list on this rank's device.  Returns (pruned tensors -- views into
    the gathered flat buffer, every layer present on every rank --, counts int64 [n_layers, 4]:
    numel, nonzero after, zeros after, nonzero before; flat buffer; layout).  gather=False skips both
    collectives (prune-only timing): then only the owned layers' views are meaningful.  gather='peer_store'
    (CUDA, world > 1): the flat buffer lives in symmetric memory and the kernels store into every rank's copy
    themselves; one device-side barrier replaces the all-gather (pass `peer_buffer` to reuse an allocation)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    layout = layout or ShardLayout([w.shape for w in weights], world)
    dev = weights[0].device if weights else torch.device('cpu')
    peer = gather == 'peer_store' and world > 1
    if peer:
        if peer_buffer is None:
            peer_buffer = PeerBuffer(layout, dev, group)
        else:
            # a reused buffer: a fast rank must not store over NVLink into a peer that is still reading the previous
            # call's result (PeerBuffer.__init__ ends with the same barrier for a fresh one)
            peer_buffer.barrier()
        flat = peer_buffer.flat
    else:
        flat = torch.zeros(world * layout.shard_elems, dtype=torch.float32, device=dev)
    mine = layout.owned(rank)
    my_shard = flat[rank * layout.shard_elems:(rank + 1) * layout.shard_elems]
    xs = [weights[i].contiguous() for i in mine]
    outs = [layout.view(flat, i) for i in mine]          # kernels store straight into the send slot
    counts_local = torch.zeros((len(weights), 4), dtype=torch.int64)
    my_levels = [int(levels[i]) for i in mine]
    c = None
    if peer:
        # A rank whose layers the mirroring kernels do not cover (generic tiled path ...) gets NotImplementedError from
        # Plan.set_peers BEFORE anything is enqueued.  A rank-local exception cannot be recovered collectively, so the
        # ranks agree first: all of them store to their peers, or all of them fall back to the NCCL all-gather together.
        ok = 1
        try:
            if mine:
                c = prune_fn(xs, outs, my_levels, peer_deltas=peer_buffer.deltas)
        except NotImplementedError:
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if not int(flag.item()):
            peer = False
            c = None       # ranks that did run redo their shard without mirroring: same values, the gather overwrites the rest
    if c is None and mine:
        c = prune_fn(xs, outs, my_levels)
    if mine:
        counts_local[mine] = torch.from_numpy(np.asarray(c, dtype=np.int64))
    counts = counts_local.to(dev)
    if peer:
        peer_buffer.barrier()
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    elif gather and world > 1:
        dist.all_gather_into_tensor(flat, my_shard.clone() if dev.type == 'cpu' else my_shard, group=group)
        # every layer has exactly one owner and the others contribute zeros: a sum is the gather of the counts
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    pruned = [layout.view(flat, i) for i in range(len(weights))]
    return pruned, counts.cpu().numpy(), flat, layout


def prune_model_sharded(weights: List[torch.Tensor], wavelet: str, level: int, percentile: float, *,
                        mode: str = 'periodization', clamp_level: bool = True, numpy_semantics: Optional[str] = None,
                        group: Optional[dist.ProcessGroup] = None, gather=True):
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
