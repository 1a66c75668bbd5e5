"""CPU checks of the C-ABI boundary: the library loads, exports every symbol include/wtprune.h
declares, and its host-only metadata entry points agree with the oracle.  No compute calls."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WAVELETS = ['haar', 'db1', 'db2', 'db4', 'db6', 'coif1', 'coif2', 'coif3', 'bior1.3', 'bior2.2', 'bior4.4',
            'rbio1.3', 'rbio2.2', 'rbio4.4', 'sym2', 'sym4', 'sym6',
            # round 2: beyond the reference's flag enum (wavelet is an open string in dwt_pruning.py:35)
            'db3', 'db5', 'db7', 'db8', 'db9', 'sym3', 'bior1.1', 'rbio1.1', 'bior3.1', 'rbio3.1']


def test_header_symbols_exported(wtlib):
    hdr = open(os.path.join(ROOT, 'include', 'wtprune.h')).read()
    declared = set(re.findall(r'\b(wt_[a-z0-9_]+)\s*\(', hdr))
    assert declared, 'no declarations parsed'
    L = wtlib.lib()
    for name in sorted(declared):
        assert hasattr(L, name), f'{name} declared in wtprune.h but not exported'
    assert declared == set(wtlib.EXPORTS)
    assert L.wt_version() == 1


def test_wavelet_tables_match_oracle(wtlib, oracle_lib):
    for w in WAVELETS:
        wid, F = wtlib.wavelet_lookup(w)
        assert F == oracle_lib.dec_len(w)
        for a, b in zip(wtlib.wavelet_filters(w), oracle_lib.filters(w)):
            np.testing.assert_array_equal(a, b)
    with pytest.raises(ValueError):
        wtlib.wavelet_lookup('nope')


def test_max_level_and_shapes(wtlib, oracle_lib):
    for w in WAVELETS:
        F = oracle_lib.dec_len(w)
        for n in range(1, 80):
            assert wtlib.dwt_max_level(n, F) == oracle_lib.dwt_max_level(n, F)
        for mode in ('periodization', 'symmetric', 'zero', 'smooth', 'antireflect'):
            for (H, W) in [(1, 1), (3, 3), (7, 7), (5, 12), (64, 64)]:
                for L in range(0, 6):
                    sh, sw, nc = wtlib.packed_shape(H, W, w, L, mode)
                    assert (sh, sw) == oracle_lib.packed_shape((H, W), w, L, mode)
                    assert nc == oracle_lib.num_coeffs((H, W), w, L, mode)
    # n in {1,3,7}: F=2 -> {0,1,2}; F=4 -> {0,0,1}; F>=6 -> 0
    assert [wtlib.dwt_max_level(n, 2) for n in (1, 3, 7)] == [0, 1, 2]
    assert [wtlib.dwt_max_level(n, 4) for n in (1, 3, 7)] == [0, 0, 1]
    assert [wtlib.dwt_max_level(n, 6) for n in (1, 3, 7)] == [0, 0, 0]


def test_percentile_plan_matches_oracle_and_numpy(wtlib, oracle_lib):
    rng = np.random.default_rng(0)
    for _ in range(300):
        n = int(rng.integers(1, 1 << 26))
        p = float(rng.choice([0, 10, 23.6, 38.2, 50, 61.8, 78.6, 90, 100, rng.uniform(0, 100)]))
        for sem, si in (('legacy_1_22', 0), ('nep50_2x', 1)):
            assert wtlib.percentile_plan(n, p, sem) == oracle_lib.percentile_plan(n, p, si)
    # live numpy (NEP 50) index arithmetic
    for n in (1, 2, 9408, 2359296, (1 << 24) + 3):
        for p in (0, 10, 50, 90, 99.99999, 100):
            q = np.true_divide(p, np.float32(100))
            vi = (n - 1) * q
            k0, k1, g = wtlib.percentile_plan(n, p, 'nep50_2x')
            if vi >= np.float32(n - 1):
                assert (k0, k1) == (n - 1, n - 1)
            else:
                assert k0 == int(np.floor(vi)) and g == float(vi - np.floor(vi))
    with pytest.raises(ValueError):
        wtlib.percentile_plan(10, 101, 'legacy_1_22')


def test_no_cuda_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip('CUDA present')
    from wavelettransforms_b200 import dwt_pruning
    with pytest.raises(RuntimeError):
        dwt_pruning.multi_resolution_analysis([torch.zeros(2, 2, 3, 3)], 'haar', 1, 50)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'wavelettransforms_b200')
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(dp, fn)).read()
                assert not re.search(r'^\s*(import|from)\s+oracle', src, re.M), fn
                assert 'oracle/' not in src.replace('oracle/dwt_oracle.c', '') or fn.endswith(('.cu', '.cuh')), fn
