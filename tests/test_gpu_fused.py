"""The fused recompute kernels (wt_fused.cuh) against the oracle, bit for bit, on both threshold kinds.  The library routes
percentile plans to the round-1 kernels by default (measured faster); WT_FUSED=all forces the fused kernels so that
path stays covered.  Reference flows: ResNet/dwt_pruning.py:35-95, ResNet/dwt_pruning_NoEntropy.py:12-62."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _weights(seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(64, 3, 7, 7, generator=g) * 0.05, torch.randn(96, 64, 3, 3, generator=g) * 0.02,
            torch.randn(40, 33, 3, 3, generator=g) * 0.03, torch.randn(128, 64, 1, 1, generator=g) * 0.02,
            torch.randn(17, 5, 5, 5, generator=g) * 0.02, torch.randn(1000, generator=g)]


def _bits(t):
    return t.detach().cpu().contiguous().numpy().view(np.uint32)


@pytest.mark.parametrize('mode', ['all', 'abs', 'off'])
@pytest.mark.parametrize('wavelet,level,clamp,pct', [('haar', 5, True, 50.0), ('db2', 2, False, 90.0), ('bior4.4', 5, True, 37.5),
                                                     ('bior2.2', 3, False, 70.0), ('haar', 1, False, 0.0), ('db2', 1, False, 100.0)])
def test_percentile_flow_bitwise(monkeypatch, mode, wavelet, level, clamp, pct):
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import dwt_pruning
    monkeypatch.setenv('WT_FUSED', mode)
    ws = _weights()
    got, gz = dwt_pruning.multi_resolution_analysis(ws, wavelet, level, pct, clamp_level=clamp)
    want, wz = rf.multi_resolution_analysis(ws, wavelet, level, pct, numpy_semantics='legacy_1_22', clamp_level=clamp)
    assert gz == wz
    for a, b in zip(got, want):
        assert np.array_equal(_bits(a), _bits(b))


@pytest.mark.parametrize('mode', ['all', 'abs', 'off'])
@pytest.mark.parametrize('wavelet,level,thr', [('haar', 2, 0.02), ('db2', 1, 0.01), ('bior2.2', 3, 0.02), ('sym4', 0, 0.015)])
def test_absolute_flow_bitwise(monkeypatch, mode, wavelet, level, thr):
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import dwt_pruning_NoEntropy
    monkeypatch.setenv('WT_FUSED', mode)
    ws = _weights(1)
    got, gp = dwt_pruning_NoEntropy.multi_resolution_analysis(ws, wavelet, level, thr)
    want, wp = rf.multi_resolution_analysis_absolute(ws, wavelet, level, thr)
    assert gp == wp
    for a, b in zip(got, want):
        assert np.array_equal(_bits(a), _bits(b))


def test_heavy_ties_take_the_exact_path(monkeypatch):
    """quantised weights: the sampled bracket cannot separate the ranks, the exact radix passes must (level 0 and haar)"""
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import dwt_pruning
    monkeypatch.setenv('WT_FUSED', 'all')
    g = torch.Generator().manual_seed(3)
    ws = [(torch.randint(-3, 4, (256, 64, 3, 3), generator=g).float() * 0.25), torch.zeros(64, 64, 3, 3),
          (torch.randint(0, 2, (300, 300), generator=g).float())]
    for wavelet, clamp in (('bior4.4', True), ('haar', True)):
        for pct in (10.0, 50.0, 93.0):
            got, gz = dwt_pruning.multi_resolution_analysis(ws, wavelet, 5, pct, clamp_level=clamp)
            want, wz = rf.multi_resolution_analysis(ws, wavelet, 5, pct, numpy_semantics='legacy_1_22', clamp_level=clamp)
            assert gz == wz, (wavelet, pct)
            for a, b in zip(got, want):
                assert np.array_equal(_bits(a), _bits(b)), (wavelet, pct)


def test_repeated_runs_reuse_the_scratch(monkeypatch):
    """the kernel of run n zeroes the scratch of run n + 1: three runs of one plan must agree"""
    from wavelettransforms_b200 import _lib, engine
    monkeypatch.setenv('WT_FUSED', 'all')
    dev = torch.device('cuda', 0)
    xs = [w.to(dev) for w in _weights(2)[:4]]
    outs = [torch.empty_like(x) for x in xs]
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (1,) for x, o in zip(xs, outs)]
    plan = _lib.Plan(items, 'haar', 'periodization', _lib.THRESHOLD_PERCENTILE, 60.0, 'legacy_1_22')
    s = int(torch.cuda.current_stream().cuda_stream)
    ref = None
    for _ in range(3):
        plan.run(s)
        c, nzb, thr = plan.results(s)
        cur = (c.copy(), nzb.copy(), thr.copy(), [o.clone() for o in outs])
        if ref is not None:
            assert np.array_equal(ref[0], cur[0]) and np.array_equal(ref[1], cur[1]) and np.array_equal(ref[2], cur[2])
            for a, b in zip(ref[3], cur[3]):
                assert torch.equal(a, b)
        ref = cur
    plan.close()
