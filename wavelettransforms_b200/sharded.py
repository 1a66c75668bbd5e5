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
        # NVSwitch multicast mapping of the same buffer (0 when the fabric / driver has no NVLS): one multimem.st reaches every rank
        mc = int(getattr(self.handle, 'multicast_ptr', 0) or 0)
        self.multicast_delta = (mc - mine) if mc else None
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
    themselves; one device-side barrier replaces the all-gather (pass `peer_buffer` to reuse an allocation).
    gather='multicast': the same through the NVSwitch multicast mapping of the buffer (one multimem.st per value, replicated
    inside the switch: a rank sends its shard once instead of world - 1 times); peer stores where there is no NVLS."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    layout = layout or ShardLayout([w.shape for w in weights], world)
    dev = weights[0].device if weights else torch.device('cpu')
    peer = gather in ('peer_store', 'multicast') and world > 1
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
                use_mc = gather == 'multicast' and peer_buffer.multicast_delta is not None
                c = prune_fn(xs, outs, my_levels,
                             peer_deltas=('multicast', peer_buffer.multicast_delta) if use_mc else peer_buffer.deltas)
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


def layer_groups(numels: Sequence[int], n_groups: int) -> List[List[int]]:
    """Layers in descending size, cut into `n_groups` runs of about equal total size (so the first group holds the few
    largest layers): the unit of the pipelined gather.  Deterministic from the shapes: every rank computes the same."""
    order = sorted(range(len(numels)), key=lambda i: (-numels[i], i))
    total = sum(numels)
    groups, acc, cur = [], 0, []
    for i in order:
        cur.append(i)
        acc += numels[i]
        if len(groups) < n_groups - 1 and acc >= total * (len(groups) + 1) / n_groups:
            groups.append(cur)
            cur = []
    if cur:
        groups.append(cur)
    return groups


def prune_sharded_pipelined(weights: List[torch.Tensor], levels: Sequence[int], prune_fn: PruneFn, n_groups: int = 4,
                            group: Optional[dist.ProcessGroup] = None):
    """prune_sharded with the gather OVERLAPPED with the pruning: the layers are dealt into `n_groups` size-ordered groups,
    each with its own LPT assignment and flat buffer; as soon as the kernels of group g are enqueued its all-gather starts
    on a side stream while group g + 1 is pruned (round 1 gathered once, after the last kernel: at 8 GPUs the gather was
    60 % of the step with nothing to hide behind).  Same results as prune_sharded; returns (pruned tensors, counts)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    dev = weights[0].device if weights else torch.device('cpu')
    cuda = dev.type == 'cuda'
    pruned: List[Optional[torch.Tensor]] = [None] * len(weights)
    counts_local = torch.zeros((len(weights), 4), dtype=torch.int64)
    side = torch.cuda.Stream(device=dev) if cuda and world > 1 else None
    main = torch.cuda.current_stream(dev) if cuda else None
    for idx in layer_groups([int(w.numel()) for w in weights], n_groups):
        layout = ShardLayout([weights[i].shape for i in idx], world)
        flat = torch.zeros(world * layout.shard_elems, dtype=torch.float32, device=dev)
        mine = layout.owned(rank)
        if mine:
            xs = [weights[idx[k]].contiguous() for k in mine]
            outs = [layout.view(flat, k) for k in mine]
            c = prune_fn(xs, outs, [int(levels[idx[k]]) for k in mine])
            counts_local[[idx[k] for k in mine]] = torch.from_numpy(np.asarray(c, dtype=np.int64))
        if world > 1:
            my_shard = flat[rank * layout.shard_elems:(rank + 1) * layout.shard_elems]
            if side is not None:
                side.wait_stream(main)                 # this group's kernels are enqueued: its gather may start behind them
                with torch.cuda.stream(side):
                    dist.all_gather_into_tensor(flat, my_shard, group=group)
                flat.record_stream(side)
            else:
                dist.all_gather_into_tensor(flat, my_shard.clone(), group=group)
        for k, i in enumerate(idx):
            pruned[i] = layout.view(flat, k)
    if side is not None:
        main.wait_stream(side)
    counts = counts_local.to(dev)
    if world > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    return pruned, counts.cpu().numpy()


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


# ---- intra-layer split: ONE tensor's planes over the ranks -----------------------------------------
# The planes of a tensor transform independently, but its percentile is ONE order statistic over every plane's
# coefficients (dwt_pruning.py:27 on the packed array of the whole tensor).  Each rank transforms its planes, the ranks
# agree on the threshold by an exact distributed radix select -- per pass a digit histogram of the local keys, summed by an
# all-reduce of at most 32 KB -- and every rank then thresholds and reconstructs its own planes.  This is the one place on
# the path with a real exchange step; layer sharding above needs none.

HistFn = Callable[[int, int, int], np.ndarray]    # (prefix, prefix_bits, digit_bits) -> local int64 histogram [2**digit_bits]
KEY_BITS = 31
RADIX_DIGITS = (12, 10, 9)


def global_kth_key(k: int, local_hist: HistFn, group=None) -> int:
    """the k-th smallest (0-based) 31-bit key of the union of every rank's keys: three passes, one all-reduce each"""
    prefix, prefix_bits = 0, 0
    for digit_bits in RADIX_DIGITS:
        h = torch.as_tensor(np.asarray(local_hist(prefix, prefix_bits, digit_bits), dtype=np.int64))
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            dev_h = h.cuda() if dist.get_backend(group) == 'nccl' else h
            dist.all_reduce(dev_h, op=dist.ReduceOp.SUM, group=group)
            h = dev_h.cpu()
        c = np.cumsum(h.numpy())
        d = int(np.searchsorted(c, k, side='right'))
        if d >= len(c):
            raise ValueError(f'rank {k} beyond the {int(c[-1])} keys under prefix {prefix:#x}')
        k -= int(c[d - 1]) if d else 0
        prefix, prefix_bits = (prefix << digit_bits) | d, prefix_bits + digit_bits
    return prefix


def percentile_lerp(a: np.float32, b: np.float32, gamma: float, numpy_semantics: str) -> np.float32:
    """the interpolation step of np.percentile exactly as libwtprune's kernels do it (csrc/wt_select.cu: lerp_threshold):
    'legacy_1_22' = (b - a) in float32, the rest in float64, result demoted; 'nep50_2x' = all float32"""
    a, b = np.float32(a), np.float32(b)
    if numpy_semantics == 'legacy_1_22':
        d = np.float64(np.float32(b - a))
        g = np.float64(gamma)
        t = np.float64(b) - d * (np.float64(1.0) - g) if g >= 0.5 else np.float64(a) + d * g
        return np.float32(t)
    g = np.float32(gamma)
    d = np.float32(b - a)
    return np.float32(b - np.float32(d * np.float32(np.float32(1.0) - g))) if g >= np.float32(0.5) else np.float32(a + np.float32(d * g))


def plane_slice(batch: int, rank: int, world: int) -> Tuple[int, int]:
    """contiguous share of `batch` planes for `rank` (the first batch % world ranks take one more)"""
    q, r = divmod(batch, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def prune_tensor_split(w: torch.Tensor, wavelet: str, level: int, percentile: float, *, mode: str = 'periodization',
                       numpy_semantics: Optional[str] = None, group=None, gather: bool = True):
    """multi_resolution_analysis([w], ...) (dwt_pruning.py:35-95, no level clamp applied here: pass the level you want) with
    the planes of `w` split over the ranks of `group`.  Every rank passes the same full tensor (host or device; only its own
    planes are uploaded).  Returns (pruned, threshold, zeros): the whole pruned tensor on every rank when gather=True, else
    this rank's planes [lo:hi] of the flattened batch; zeros = count over the whole tensor.  Bit-identical to the
    single-GPU plan: same kernels, same order statistics, same interpolation.

    One plan per rank, run in two stages (wt_plan_run_stage): the analysis leaves the coefficients in the plan's workspace,
    the ranks agree on the threshold (three histogram passes over them + three all-reduces per order statistic), the
    threshold goes into the plan's threshold array, and the second stage thresholds and reconstructs."""
    from . import _lib, config, engine
    sem = numpy_semantics or config.numpy_semantics
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    dev = engine.cuda_device()
    batch, H, W = engine.plane_geometry(w.shape)
    if w.ndim < 2:
        level = 0
    lo, hi = plane_slice(batch, rank, world)
    sh, sw, _ = _lib.packed_shape(H, W, wavelet, level, mode)
    n_total = batch * sh * sw                                  # what np.percentile sees: the packed array, padding zeros included
    k0, k1, gamma = _lib.percentile_plan(n_total, float(percentile), sem)
    L, s = _lib.lib(), int(torch.cuda.current_stream().cuda_stream)
    plan, x, out = None, None, None
    coef_ptr, n_explicit, n_implicit = 0, 0, 0
    if hi > lo:
        x = engine.to_device_f32(w.reshape(batch, H, W)[lo:hi], dev)
        out = torch.empty_like(x)
        plan = _lib.Plan([(x.data_ptr(), out.data_ptr(), hi - lo, H, W, level)], wavelet, mode, _lib.THRESHOLD_PERCENTILE,
                         float(percentile), sem)
    try:
        if plan is not None:
            plan.run_stage(1, s)
            coef_ptr, n_explicit, n_implicit = plan.layer_coefficients(0)
        zt = torch.tensor([n_implicit], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(zt, op=dist.ReduceOp.SUM, group=group)
        implicit_total = int(zt.item())                        # padding cells of coeffs_to_array over the whole tensor
        hist = torch.zeros(1 << max(RADIX_DIGITS), dtype=torch.int64, device=dev)

        def local_hist(prefix, prefix_bits, digit_bits):
            hist.zero_()
            if n_explicit:
                _lib.check(L.wt_key_hist_f32(coef_ptr, n_explicit, prefix, prefix_bits, digit_bits, hist.data_ptr(), s))
            return hist[:1 << digit_bits].cpu().numpy()

        def kth(k):                                            # the implicit zeros sort first
            return 0 if k < implicit_total else global_kth_key(k - implicit_total, local_hist, group)

        ka = kth(k0)
        kb = ka if k1 == k0 else kth(k1)
        a, b = np.array([ka, kb], dtype=np.uint32).view(np.float32)
        t = percentile_lerp(a, b, gamma, sem)
        zeros = 0
        if plan is not None:
            _, thr_ptr = plan.device_result_pointers()
            torch.as_tensor(engine._DeviceBlock(thr_ptr, (1,), '<f4'), device=dev).fill_(float(t))
            plan.run_stage(2, s)
            counts, _, _ = plan.results(s)
            zeros = int(counts[0, 2])
    finally:
        if plan is not None:
            plan.close()
    zt = torch.tensor([zeros], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(zt, op=dist.ReduceOp.SUM, group=group)
    if out is None:
        out = torch.empty((0, H, W), dtype=torch.float32, device=dev)
    if not gather:
        return out, np.float32(t), int(zt.item())
    if world == 1:
        return out.reshape(w.shape), np.float32(t), int(zt.item())
    q = (batch + world - 1) // world                          # padded equal shares for all_gather_into_tensor
    send = torch.zeros((q, H, W), dtype=torch.float32, device=dev)
    send[:hi - lo] = out
    recv = torch.empty((world * q, H, W), dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(recv, send, group=group)
    parts = []
    for r in range(world):
        a0, a1 = plane_slice(batch, r, world)
        parts.append(recv[r * q:r * q + (a1 - a0)])
    return torch.cat(parts, 0).reshape(w.shape), np.float32(t), int(zt.item())
