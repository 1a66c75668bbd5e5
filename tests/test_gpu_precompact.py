"""WT_PRECOMPACT=1: the small-plane analysis kernels do the select's sample and count / compact passes (wt_small.cu MODE 1 / 2,
wt_select_bracket_kernel, wt_select_kernel path 2).  Opt-in because it measured slower; it has to stay
bit-exact against the oracle all the same.  Reference flow: ResNet/dwt_pruning.py:35-95."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _weights(seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(64, 3, 7, 7, generator=g) * 0.05, torch.randn(192, 128, 3, 3, generator=g) * 0.02,
            torch.randn(40, 33, 3, 3, generator=g) * 0.03, torch.randn(128, 64, 1, 1, generator=g) * 0.02,
            torch.randn(17, 5, 5, 5, generator=g) * 0.02, torch.randn(1000, generator=g)]


def _bits(t):
    return t.detach().cpu().contiguous().numpy().view(np.uint32)


@pytest.mark.parametrize('wavelet,level,clamp,pct', [('haar', 5, True, 50.0), ('db2', 2, False, 90.0), ('bior4.4', 5, False, 90.0),
                                                     ('bior2.2', 3, False, 70.0), ('haar', 1, False, 0.0), ('db2', 1, False, 100.0),
                                                     ('bior4.4', 5, True, 37.5)])
def test_precompacted_percentile_flow_bitwise(monkeypatch, wavelet, level, clamp, pct):
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import _lib, dwt_pruning, engine
    monkeypatch.setenv('WT_PRECOMPACT', '1')
    ws = _weights()
    got, gz = dwt_pruning.multi_resolution_analysis(ws, wavelet, level, pct, clamp_level=clamp)
    want, wz = rf.multi_resolution_analysis(ws, wavelet, level, pct, numpy_semantics='legacy_1_22', clamp_level=clamp)
    assert gz == wz
    for a, b in zip(got, want):
        assert np.array_equal(_bits(a), _bits(b))


def test_precompacted_plan_reports_its_path_and_repeats(monkeypatch):
    """the plan says which select it runs; three runs of one plan (alternating zero blocks) agree bit for bit"""
    from wavelettransforms_b200 import _lib, engine
    monkeypatch.setenv('WT_PRECOMPACT', '1')
    dev = torch.device('cuda', 0)
    xs = [w.to(dev) for w in _weights(2)[:3]]
    outs = [torch.empty_like(x) for x in xs]
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (1,) for x, o in zip(xs, outs)]
    plan = _lib.Plan(items, 'bior2.2', 'periodization', _lib.THRESHOLD_PERCENTILE, 80.0, 'legacy_1_22')
    try:
        assert plan.select_path == 'precompacted'
        s = int(torch.cuda.current_stream().cuda_stream)
        ref = None
        for _ in range(3):
            plan.run(s)
            torch.cuda.synchronize()
            cur = [_bits(o).copy() for o in outs]
            if ref is None:
                ref = cur
            for a, b in zip(cur, ref):
                assert np.array_equal(a, b)
    finally:
        plan.close()
    monkeypatch.delenv('WT_PRECOMPACT')
    plan = _lib.Plan(items, 'bior2.2', 'periodization', _lib.THRESHOLD_PERCENTILE, 80.0, 'legacy_1_22')
    try:
        assert plan.select_path != 'precompacted'
        plan.run(int(torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        for o, b in zip(outs, ref):
            assert np.array_equal(_bits(o), b)
    finally:
        plan.close()


def test_precompacted_heavy_ties_fall_back(monkeypatch):
    """quantised weights: the sampled bracket cannot separate the ranks; the radix passes of the same launch must"""
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import dwt_pruning
    monkeypatch.setenv('WT_PRECOMPACT', '1')
    g = torch.Generator().manual_seed(3)
    ws = [(torch.randint(-3, 4, (256, 64, 3, 3), generator=g).float() * 0.25), torch.zeros(64, 64, 3, 3),
          (torch.randint(0, 2, (300, 300), generator=g).float())]
    for wavelet, clamp in (('bior4.4', True), ('haar', True)):
        for pct in (10.0, 93.0):
            got, gz = dwt_pruning.multi_resolution_analysis(ws, wavelet, 5, pct, clamp_level=clamp)
            want, wz = rf.multi_resolution_analysis(ws, wavelet, 5, pct, numpy_semantics='legacy_1_22', clamp_level=clamp)
            assert gz == wz, (wavelet, pct)
            for a, b in zip(got, want):
                assert np.array_equal(_bits(a), _bits(b)), (wavelet, pct)
