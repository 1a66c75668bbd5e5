"""The named fast mode `fma_taporder` (wt_plan_set_arithmetic): every DWT multiply-add fused into one FFMA, taps
in the oracle's order.  Bit-exact against the oracle's twin (orc_set_arithmetic(1)), within the north star's tolerance
(1e-5 relative / 1e-6 absolute) of the faithful weights; never the default.  Reference call sites:
ResNet/dwt_pruning.py:67-77, ResNet/dwt_pruning_NoEntropy.py:41-49."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _weights(seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(64, 3, 7, 7, generator=g) * 0.05, torch.randn(96, 64, 3, 3, generator=g) * 0.02,
            torch.randn(128, 64, 1, 1, generator=g) * 0.02, torch.randn(17, 5, 5, 5, generator=g) * 0.02]


def _bits(t):
    return t.detach().cpu().contiguous().numpy().view(np.uint32)


@pytest.mark.parametrize('wavelet,level,clamp,pct', [('bior4.4', 5, False, 90.0), ('bior2.2', 3, False, 50.0), ('haar', 5, True, 70.0),
                                                     ('db4', 2, False, 80.0), ('coif3', 1, False, 60.0), ('rbio1.3', 4, False, 95.0)])
def test_fma_taporder_is_bitwise_the_oracle_twin(oracle_lib, wavelet, level, clamp, pct):
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import dwt_pruning
    ws = _weights()
    got, gz = dwt_pruning.multi_resolution_analysis(ws, wavelet, level, pct, clamp_level=clamp, arithmetic='fma_taporder')
    with oracle_lib.arithmetic('fma_taporder'):
        want, wz = rf.multi_resolution_analysis(ws, wavelet, level, pct, numpy_semantics='legacy_1_22', clamp_level=clamp)
    assert gz == wz
    for a, b in zip(got, want):
        assert np.array_equal(_bits(a), _bits(b))
    # and close to the faithful arithmetic wherever the same coefficients survive the threshold: absolute flow, no select
    from wavelettransforms_b200 import dwt_pruning_NoEntropy as ne
    fast, _ = ne.multi_resolution_analysis(ws, wavelet, level, 0.0, arithmetic='fma_taporder')
    slow, _ = ne.multi_resolution_analysis(ws, wavelet, level, 0.0)
    worst = 0.0
    for a, b in zip(fast, slow):
        np.testing.assert_allclose(a.numpy(), b.numpy(), rtol=1e-5, atol=1e-6)
        worst = max(worst, float((a - b).abs().max()))
    print(f'{wavelet} L{level}: max |fma_taporder - faithful| over the round trip = {worst:.3e}')


def test_fma_taporder_absolute_flow_and_default_stays_faithful(oracle_lib):
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import config, dwt_pruning_NoEntropy as ne
    assert config.arithmetic == 'faithful'
    ws = _weights(1)
    got, gp = ne.multi_resolution_analysis(ws, 'bior4.4', 5, 0.02, arithmetic='fma_taporder')
    with oracle_lib.arithmetic('fma_taporder'):
        want, wp = rf.multi_resolution_analysis_absolute(ws, 'bior4.4', 5, 0.02)
    assert gp == wp
    for a, b in zip(got, want):
        assert np.array_equal(_bits(a), _bits(b))
    got, gp = ne.multi_resolution_analysis(ws, 'bior4.4', 5, 0.02)           # default: the faithful arithmetic
    want, wp = rf.multi_resolution_analysis_absolute(ws, 'bior4.4', 5, 0.02)
    assert gp == wp and all(np.array_equal(_bits(a), _bits(b)) for a, b in zip(got, want))


def test_unknown_arithmetic_is_refused_and_generic_plans_take_the_fast_mode():
    from wavelettransforms_b200 import _lib, engine
    dev = torch.device('cuda', 0)
    x = torch.randn(2, 64, 64, device=dev)
    o = torch.empty_like(x)
    plan = _lib.Plan([(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (2,)], 'db4', 'periodization',
                     _lib.THRESHOLD_PERCENTILE, 50.0, 'legacy_1_22')
    plan.set_arithmetic('fma_taporder')      # round 2: the tiled kernels exist in both arithmetics (wt_generic_fma.cu)
    with pytest.raises(ValueError):
        plan.set_arithmetic('fast')
    plan.close()


@pytest.mark.parametrize('mode', ['periodization', 'symmetric', 'reflect'])
@pytest.mark.parametrize('wavelet,level', [('db4', 3), ('bior4.4', 2), ('haar', 2), ('db8', 1)])
def test_fma_taporder_on_the_large_plane_and_generic_kernels(oracle_lib, wavelet, level, mode):
    """the tiled kernels (wt_generic_fma.cu): TMA-staged fixed-geometry kernels on the big planes, runtime-geometry kernels on
    the odd one, two-sweep kernels for the extra mode -- both threshold kinds, bit for bit against the oracle's twin"""
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import dwt_pruning, dwt_pruning_NoEntropy as ne
    g = torch.Generator().manual_seed(31)
    ws = [torch.randn(1, 1024, 1280, generator=g) * 0.02, torch.randn(2, 530, 517, generator=g) * 0.02, torch.randn(3, 2, 9, 12, generator=g) * 0.02]
    got, gz = dwt_pruning.multi_resolution_analysis(ws, wavelet, level, 83.0, mode, clamp_level=False, arithmetic='fma_taporder')
    with oracle_lib.arithmetic('fma_taporder'):
        want, wz = rf.multi_resolution_analysis(ws, wavelet, level, 83.0, mode=mode, numpy_semantics='legacy_1_22', clamp_level=False)
    assert gz == wz
    for a, b in zip(got, want):
        assert np.array_equal(_bits(a), _bits(b))
    got, gp = ne.multi_resolution_analysis(ws, wavelet, level, 0.015, mode, arithmetic='fma_taporder')
    with oracle_lib.arithmetic('fma_taporder'):
        want, wp = rf.multi_resolution_analysis_absolute(ws, wavelet, level, 0.015, mode)
    assert gp == wp
    for a, b in zip(got, want):
        assert np.array_equal(_bits(a), _bits(b))
    # the two arithmetics do differ on these inputs (a silent fall-back to the faithful kernels would pass the checks above
    # only if the oracle twin were broken too, but make it explicit)
    slow, _ = ne.multi_resolution_analysis(ws, wavelet, level, 0.015, mode)
    assert any(not np.array_equal(_bits(a), _bits(b)) for a, b in zip(got, slow)) or wavelet == 'haar'
    for a, b in zip(got, slow):
        np.testing.assert_allclose(a.numpy(), b.numpy(), rtol=1e-5, atol=1e-6)
