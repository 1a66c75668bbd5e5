"""wt_plan_set_graph: a plan replayed as a CUDA graph (one graph per alternating scratch block) gives the plain run's bits,
counts and thresholds -- on the percentile select, the absolute-threshold kernels, the fused level-0 kernel and the generic
tiled path.  The oracle pins the plain run elsewhere (tests/test_gpu_parity.py); here the plain run is the yardstick."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bits(t):
    return t.detach().cpu().contiguous().numpy().view(np.uint32)


def _weights(seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(64, 3, 7, 7, generator=g) * 0.05, torch.randn(96, 64, 3, 3, generator=g) * 0.02,
            torch.randn(128, 64, 1, 1, generator=g) * 0.02, torch.randn(2, 1, 200, 264, generator=g) * 0.1]


def _run(xs, levels, wavelet, kind, value, graph, runs, arithmetic=None):
    from wavelettransforms_b200 import _lib, engine
    outs = [torch.empty_like(x) for x in xs]
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,) for x, o, lv in zip(xs, outs, levels)]
    plan = _lib.Plan(items, wavelet, 'periodization', kind, value, 'legacy_1_22')
    s = torch.cuda.Stream()
    res = []
    try:
        if graph is not None:
            plan.set_graph(graph)   # None: the library's default (automatic for large-plane plans)
        with torch.cuda.stream(s):
            for r in range(runs):
                if arithmetic and r == runs - 2:
                    plan.set_arithmetic(arithmetic)   # drops the graphs: the next runs re-capture
                for o in outs:
                    o.fill_(float('nan'))
                plan.run(int(s.cuda_stream))
                counts, nzb, thr = plan.results(int(s.cuda_stream))
                res.append(([_bits(o).copy() for o in outs], np.array(counts), np.array(nzb), np.array(thr).view(np.uint32)))
        launches = plan.num_launches
    finally:
        plan.close()
    return res, launches


@pytest.mark.parametrize('wavelet,levels,kind,value', [
    ('bior2.2', (2, 1, 0, 3), 'percentile', 80.0),
    ('haar', (2, 1, 0, 0), 'percentile', 55.0),
    ('db2', (1, 1, 0, 2), 'absolute', 0.02),
    ('haar', (0, 0, 0, 0), 'absolute', 0.015),
])
def test_graph_replay_equals_plain_runs(wavelet, levels, kind, value):
    from wavelettransforms_b200 import _lib
    dev = torch.device('cuda', 0)
    xs = [w.to(dev) for w in _weights()]
    k = _lib.THRESHOLD_PERCENTILE if kind == 'percentile' else _lib.THRESHOLD_ABSOLUTE
    plain, n_plain = _run(xs, levels, wavelet, k, value, False, 1)
    graph, n_graph = _run(xs, levels, wavelet, k, value, True, 5)   # runs 0, 1 capture, 2..4 replay
    assert n_graph == n_plain
    for r, (bits, counts, nzb, thr) in enumerate(graph):
        for a, b in zip(bits, plain[0][0]):
            assert np.array_equal(a, b), r
        assert np.array_equal(counts, plain[0][1]) and np.array_equal(nzb, plain[0][2]) and np.array_equal(thr, plain[0][3]), r


def test_graph_is_dropped_when_the_arithmetic_changes():
    from wavelettransforms_b200 import _lib
    dev = torch.device('cuda', 0)
    xs = [w.to(dev) for w in _weights(1)[:3]]
    k = _lib.THRESHOLD_PERCENTILE
    want, _ = _run(xs, (2, 1, 0), 'bior4.4', k, 70.0, False, 6, arithmetic='fma_taporder')
    got, _ = _run(xs, (2, 1, 0), 'bior4.4', k, 70.0, True, 6, arithmetic='fma_taporder')
    for (gb, gc, gn, gt), (wb, wc, wn, wt) in zip(got, want):
        for a, b in zip(gb, wb):
            assert np.array_equal(a, b)
        assert np.array_equal(gc, wc) and np.array_equal(gt, wt)
    # the two arithmetics differ somewhere on these inputs, so a stale graph would have been caught
    assert any(not np.array_equal(a, b) for a, b in zip(got[0][0], got[-1][0]))


def test_graph_on_the_legacy_default_stream_falls_back():
    """stream 0 cannot be captured: the plan runs plain launches instead of failing"""
    from wavelettransforms_b200 import _lib, engine
    dev = torch.device('cuda', 0)
    xs = [w.to(dev) for w in _weights(2)[:2]]
    outs = [torch.empty_like(x) for x in xs]
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (1,) for x, o in zip(xs, outs)]
    plan = _lib.Plan(items, 'haar', 'periodization', _lib.THRESHOLD_PERCENTILE, 50.0, 'legacy_1_22')
    try:
        plan.set_graph(True)
        torch.cuda.synchronize()
        for _ in range(3):
            plan.run(0)
        torch.cuda.synchronize()
        first = [_bits(o).copy() for o in outs]
        plan.set_graph(False)
        plan.run(0)
        torch.cuda.synchronize()
        for a, o in zip(first, outs):
            assert np.array_equal(a, _bits(o))
    finally:
        plan.close()


@pytest.mark.parametrize('kind,value', [('percentile', 65.0), ('absolute', 0.02)])
def test_device_results_match_host_results(kind, value):
    """engine.prune_device_tensors(device_results=True): counts and thresholds read from the plan's device blocks
    (wt_plan_counts_device / wt_plan_thresholds_device) equal what wt_plan_results copies to the host"""
    from wavelettransforms_b200 import _lib, engine
    dev = torch.device('cuda', 0)
    xs = [w.to(dev) for w in _weights(4)]
    k = _lib.THRESHOLD_PERCENTILE if kind == 'percentile' else _lib.THRESHOLD_ABSOLUTE
    levels = (2, 1, 0, 2)
    host = engine.prune_device_tensors(xs, levels, 'db2', 'periodization', k, value)
    devr = engine.prune_device_tensors(xs, levels, 'db2', 'periodization', k, value, device_results=True)
    assert devr.counts_dev.is_cuda and devr.thresholds_dev.is_cuda and tuple(devr.counts_dev.shape) == (len(xs), 3)
    back = devr.host()
    assert np.array_equal(back.counts, host.counts)
    assert np.array_equal(back.nonzero_before, host.nonzero_before)
    assert np.array_equal(back.thresholds.view(np.uint32), host.thresholds.view(np.uint32))
    for a, b in zip(devr.outputs, host.outputs):
        assert np.array_equal(_bits(a), _bits(b))


def test_large_plane_plans_replay_by_default():
    """no switch touched: a plan with layers on the large-plane kernels captures its launch sequence on runs 2 and 3 and
    replays it afterwards; every run gives the bits, counts and thresholds of the first (plain) one"""
    from wavelettransforms_b200 import _lib
    dev = torch.device('cuda', 0)
    g = torch.Generator().manual_seed(8)
    xs = [(torch.randn(1, 600, 520, generator=g) * 0.05).to(dev), (torch.randn(64, 32, 3, 3, generator=g) * 0.05).to(dev)]
    res, launches = _run(xs, (4, 1), 'db4', _lib.THRESHOLD_PERCENTILE, 75.0, None, 6)
    assert launches >= 8
    for r, (bits, counts, nzb, thr) in enumerate(res[1:], 1):
        for a, b in zip(bits, res[0][0]):
            assert np.array_equal(a, b), r
        assert np.array_equal(counts, res[0][1]) and np.array_equal(nzb, res[0][2]) and np.array_equal(thr, res[0][3]), r


def test_staged_runs_equal_the_whole_run_and_can_be_mixed_with_it():
    """wt_plan_run_stage: analysis and (threshold + synthesis) as two calls, the threshold taken from the plan's threshold
    array -- here the one the whole run just left there -- give the whole run's bits and counts; whole runs before and
    after a staged one (alternating scratch blocks, graph replay) stay right; the coefficient workspace is exposed"""
    from wavelettransforms_b200 import _lib, engine
    dev = torch.device('cuda', 0)
    xs = [w.to(dev) for w in _weights(6)]
    outs = [torch.empty_like(x) for x in xs]
    levels = (2, 1, 0, 2)
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,) for x, o, lv in zip(xs, outs, levels)]
    plan = _lib.Plan(items, 'bior2.2', 'periodization', _lib.THRESHOLD_PERCENTILE, 72.0, 'legacy_1_22')
    st = torch.cuda.Stream()
    try:
        with torch.cuda.stream(st):
            s = int(st.cuda_stream)

            def snapshot():
                counts, nzb, thr = plan.results(s)
                return [_bits(o).copy() for o in outs], np.array(counts), np.array(nzb), np.array(thr).view(np.uint32)

            plan.run(s)
            ref = snapshot()
            for rounds in range(2):
                for o in outs:
                    o.fill_(float('nan'))
                plan.run_stage(1, s)
                ptr, n_explicit, n_implicit = plan.layer_coefficients(0)
                b, h, w = engine.plane_geometry(xs[0].shape)
                sh, sw, ncoef = _lib.packed_shape(h, w, 'bior2.2', levels[0], 'periodization')
                assert ptr != 0 and n_explicit == b * ncoef and n_implicit == b * (sh * sw - ncoef)
                ptr0, n0, z0 = plan.layer_coefficients(2)          # level 0: the input itself
                assert ptr0 == xs[2].data_ptr() and n0 == xs[2].numel() and z0 == 0
                plan.run_stage(2, s)
                got = snapshot()
                for a, bb in zip(got[0], ref[0]):
                    assert np.array_equal(a, bb)
                assert all(np.array_equal(a, bb) for a, bb in zip(got[1:], ref[1:]))
                for _ in range(3):                                  # whole runs again: plain, capture, replay
                    for o in outs:
                        o.fill_(float('nan'))
                    plan.run(s)
                    got = snapshot()
                    for a, bb in zip(got[0], ref[0]):
                        assert np.array_equal(a, bb)
                    assert all(np.array_equal(a, bb) for a, bb in zip(got[1:], ref[1:]))
    finally:
        plan.close()
