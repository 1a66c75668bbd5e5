"""Layer-sharded pruning (wavelettransforms_b200/sharded.py): host logic on CPU with gloo, world_size 2 and 3,
the oracle's reference flow standing in for the kernels.  The GPU twin is tests/test_gpu_sharded.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


SHAPES = [(8, 3, 7, 7), (16, 8, 3, 3), (16, 16, 3, 3), (32, 16, 1, 1), (32, 16, 3, 3), (32, 32, 3, 3), (5, 3, 3, 3)]


def _weights():
    g = torch.Generator().manual_seed(7)
    return [torch.randn(s, generator=g) * 0.05 for s in SHAPES]


def _oracle_fn(wavelet, pct, clamp):
    from oracle import reference_flow as rf

    def fn(xs, outs, levels):
        c = np.zeros((len(xs), 4), dtype=np.int64)
        for i, (x, o, lv) in enumerate(zip(xs, outs, levels)):
            # per-layer call, the level is already clamped by the caller
            res, zeros = rf.multi_resolution_analysis([x], wavelet, lv, pct, numpy_semantics='legacy_1_22', clamp_level=False)
            o.copy_(res[0])
            c[i] = (x.numel(), x.numel() - zeros, zeros, int(torch.count_nonzero(x)))
        return c
    return fn


def _worker(rank, world, port, wavelet, level, pct, clamp, out_q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from oracle import reference_flow as rf
        from wavelettransforms_b200 import sharded
        ws = _weights()
        levels = [min(level, rf.calculate_max_level(w.shape, wavelet)) if clamp else level for w in ws]
        pruned, counts, flat, layout = sharded.prune_sharded(ws, levels, _oracle_fn(wavelet, pct, clamp))
        # the pipelined variant (groups of layers, one gather per group) must give the same tensors and counts
        pruned2, counts2 = sharded.prune_sharded_pipelined(ws, levels, _oracle_fn(wavelet, pct, clamp), n_groups=3)
        assert np.array_equal(counts, counts2)
        for a, b in zip(pruned, pruned2):
            assert torch.equal(a, b)
        out_q.put((rank, [p.clone().numpy() for p in pruned], counts, layout.owner))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
@pytest.mark.parametrize('wavelet,level,clamp', [('haar', 2, True), ('bior2.2', 2, False)])
def test_sharded_equals_single_process(oracle_lib, world, wavelet, level, clamp):
    from oracle import reference_flow as rf
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, wavelet, level, 50.0, clamp, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ws = _weights()
    want, want_counts = [], []
    for w in ws:
        lv = min(level, rf.calculate_max_level(w.shape, wavelet)) if clamp else level
        res, zeros = rf.multi_resolution_analysis([w], wavelet, lv, 50.0, numpy_semantics='legacy_1_22', clamp_level=False)
        want.append(res[0].numpy())
        want_counts.append((w.numel(), w.numel() - zeros, zeros, int(torch.count_nonzero(w))))
    owners = results[0][3]
    assert set(owners) == set(range(world))          # every rank got work
    for rank, pruned, counts, owner in results:
        assert owner == owners                        # same deterministic assignment on every rank
        assert np.array_equal(counts, np.array(want_counts, dtype=np.int64))
        for got, exp in zip(pruned, want):            # every layer present on every rank after the all-gather
            assert np.array_equal(got.view(np.uint32), exp.view(np.uint32))


def test_layout_and_lpt():
    from wavelettransforms_b200 import sharded, workloads
    census = workloads.conv_census('resnet18')
    shapes = [s for _, s in census]
    for world in (1, 2, 4, 8):
        lay = sharded.ShardLayout(shapes, world)
        assert sorted(sum((lay.owned(r) for r in range(world)), [])) == list(range(len(shapes)))
        # slots of one shard do not overlap and stay inside the shard
        for r in range(world):
            spans = sorted((lay.offset[i], lay.offset[i] + lay.numels[i]) for i in lay.owned(r))
            for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
                assert a1 <= b0
            assert all(e <= lay.shard_elems for _, e in spans)
            assert all(s % 64 == 0 for s, _ in spans)
        assert max(lay.load) >= sum(lay.numels) / world
    # ResNet-18's largest layer is 21 % of the model => 8-way balance ~0.59
    lay8 = sharded.ShardLayout(shapes, 8)
    assert max(lay8.load) == 2359296 and abs(lay8.balance - 0.5917) < 0.01
    assert workloads.lpt_assignment([5, 5, 4, 3, 3], 2) == [0, 1, 0, 1, 1]
    groups = sharded.layer_groups(lay8.numels, 4)
    assert sorted(sum(groups, [])) == list(range(len(shapes))) and len(groups) == 4
    assert lay8.numels[groups[0][0]] == max(lay8.numels)      # large layers first


class _FakePeerBuffer:
    """stands in for sharded.PeerBuffer on CPU: a plain flat buffer, no peers to store into, gloo barrier"""

    def __init__(self, layout, world):
        self.flat = torch.zeros(world * layout.shard_elems, dtype=torch.float32)
        self.deltas = [0] * (world - 1)
        self.barriers = 0

    def barrier(self):
        self.barriers += 1
        dist.barrier()


def _fallback_worker(rank, world, port, out_q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from wavelettransforms_b200 import sharded
        ws = _weights()
        levels = [0] * len(ws)
        base = _oracle_fn('haar', 50.0, True)
        calls = []

        def fn(xs, outs, levels, peer_deltas=None):
            calls.append(peer_deltas is not None)
            if peer_deltas is not None and rank == 1:      # this rank's layers are "off the mirrored paths"
                raise NotImplementedError('layer runs on the generic tiled path')
            return base(xs, outs, levels)
        layout = sharded.ShardLayout([w.shape for w in ws], world)
        pb = _FakePeerBuffer(layout, world)
        pruned, counts, flat, _ = sharded.prune_sharded(ws, levels, fn, layout=layout, gather='peer_store', peer_buffer=pb)
        out_q.put((rank, [p.clone().numpy() for p in pruned], counts, calls, pb.barriers))
    finally:
        dist.destroy_process_group()


def test_peer_store_falls_back_on_all_ranks_together(oracle_lib):
    """one rank refusing the peer-store mode must not leave the others blocked in the barrier: every rank
    agrees (all-reduce MIN) and all of them take the all-gather"""
    from oracle import reference_flow as rf
    world = 2
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_fallback_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = [rf.multi_resolution_analysis([w], 'haar', 0, 50.0, numpy_semantics='legacy_1_22', clamp_level=False)[0][0].numpy()
            for w in _weights()]
    for rank, pruned, counts, calls, barriers in results:
        assert calls == [True, False]          # tried with peers, then every rank redid its shard without
        assert barriers == 1                   # the entry barrier of a reused buffer only: no peer-store barrier after the fallback
        for got, exp in zip(pruned, want):
            assert np.array_equal(got.view(np.uint32), exp.view(np.uint32))


# ---- intra-layer split: the distributed exact select (host logic on CPU, numpy standing in for wt_key_hist_f32) ----------
def _kth_worker(rank, world, port, out_q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from wavelettransforms_b200 import sharded
        rng = np.random.default_rng(100)
        full = (rng.standard_normal(20011) * 0.05).astype(np.float32)
        full[::7] = 0.0                       # a massive tie at zero
        full[5::97] = full[3]                 # and ties elsewhere
        lo, hi = sharded.plane_slice(len(full), rank, world)
        keys = full[lo:hi].view(np.uint32) & np.uint32(0x7fffffff)

        def local_hist(prefix, prefix_bits, digit_bits):
            sel = keys if prefix_bits == 0 else keys[(keys >> np.uint32(31 - prefix_bits)) == np.uint32(prefix)]
            digit = (sel >> np.uint32(31 - prefix_bits - digit_bits)) & np.uint32((1 << digit_bits) - 1)
            return np.bincount(digit.astype(np.int64), minlength=1 << digit_bits)

        got = [sharded.global_kth_key(k, local_hist) for k in (0, 1, 2858, 2859, 10005, 20010)]
        out_q.put((rank, got))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
def test_distributed_kth_key_is_exact(world):
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_kth_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(100)
    full = (rng.standard_normal(20011) * 0.05).astype(np.float32)
    full[::7] = 0.0
    full[5::97] = full[3]
    s = np.sort(np.abs(full))
    want = [int(np.float32(s[k]).view(np.uint32)) for k in (0, 1, 2858, 2859, 10005, 20010)]
    for rank, got in results:
        assert got == want, rank


def test_plane_slice_and_percentile_lerp_match_the_oracle(oracle_lib):
    from wavelettransforms_b200 import sharded
    for batch in (0, 1, 5, 8, 17):
        for world in (1, 2, 3, 8):
            cuts = [sharded.plane_slice(batch, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == batch
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            assert max(h - l for l, h in cuts) - min(h - l for l, h in cuts) <= 1
    rng = np.random.default_rng(3)
    for _ in range(2000):
        a, b = np.sort(np.abs(rng.standard_normal(2).astype(np.float32) * np.float32(10.0 ** rng.integers(-20, 3))))
        g = float(rng.random()) if rng.random() < 0.9 else float(rng.integers(0, 2))
        for sem, idx in (('legacy_1_22', 0), ('nep50_2x', 1)):
            got = sharded.percentile_lerp(a, b, g, sem)
            want = oracle_lib.percentile_lerp(a, b, g, idx)
            assert np.float32(got).view(np.uint32) == np.float32(want).view(np.uint32), (a, b, g, sem)
