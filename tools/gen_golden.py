#!/usr/bin/env python
"""Writes tests/golden/oracle_golden.json: small seeded inputs with the oracle's packed coefficient
arrays and pruned outputs (hex of the float32 bytes).  PyWavelets cannot be imported in this image
(SURVEY F2), so these vectors come from oracle/ itself: they pin the oracle against drift between
machines/compilers (this container vs the GPU box) and give the CUDA parity tests fixed answers;
they do NOT pin the oracle against real pywt (parity stays "unpinned", see oracle/dwt_oracle.c)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle as o  # noqa: E402
from oracle import reference_flow as rf  # noqa: E402

CASES = [
    # name, shape, wavelet, level, mode, percentile, semantics, clamp
    ('haar_3x3_L1_clamped', (4, 3, 3, 3), 'haar', 1, 'periodization', 50.0, 'legacy_1_22', True),
    ('haar_7x7_L5_clamped', (2, 3, 7, 7), 'haar', 5, 'periodization', 50.0, 'legacy_1_22', True),
    ('db2_7x7_L5_clamped', (2, 3, 7, 7), 'db2', 5, 'periodization', 38.2, 'legacy_1_22', True),
    ('bior13_3x3_L1_clamped_level0', (4, 4, 3, 3), 'bior1.3', 1, 'periodization', 50.0, 'legacy_1_22', True),
    ('bior44_3x3_L5_clamped_level0', (4, 4, 3, 3), 'bior4.4', 5, 'periodization', 61.8, 'nep50_2x', True),
]
for wl in ('bior1.3', 'bior2.2', 'bior4.4', 'rbio1.3', 'rbio2.2', 'rbio4.4', 'db4', 'sym6', 'coif3'):
    CASES.append((f'{wl}_3x3_L5_unclamped', (3, 4, 3, 3), wl, 5, 'periodization', 50.0, 'legacy_1_22', False))
for wl in ('bior2.2', 'rbio4.4', 'db6', 'coif2'):
    CASES.append((f'{wl}_7x7_L3_unclamped', (2, 2, 7, 7), wl, 3, 'periodization', 23.6, 'nep50_2x', False))
for wl in ('haar', 'bior4.4', 'db4'):
    CASES.append((f'{wl}_1x1_L2_unclamped', (5, 7, 1, 1), wl, 2, 'periodization', 78.6, 'legacy_1_22', False))
for wl in ('db4', 'bior2.2', 'sym4'):
    CASES.append((f'{wl}_16x16_L2_sym', (2, 16, 16), wl, 2, 'symmetric', 90.0, 'legacy_1_22', False))
    CASES.append((f'{wl}_9x12_L2_per', (2, 9, 12), wl, 2, 'periodization', 10.0, 'nep50_2x', False))


def main():
    out = []
    for i, (name, shape, wl, lv, mode, p, sem, clamp) in enumerate(CASES):
        rng = np.random.default_rng(1000 + i)
        x = (rng.standard_normal(shape) * 0.02).astype(np.float32)
        eff = lv
        if clamp:
            eff = min(lv, rf.calculate_max_level(shape, wl))
        packed = o.wavedec2_packed(x, wl, eff, mode)
        pruned, zeros = rf.multi_resolution_analysis([torch.from_numpy(x)], wl, lv, p, mode=mode, numpy_semantics=sem,
                                                     clamp_level=clamp)
        out.append(dict(name=name, shape=list(shape), wavelet=wl, level=lv, effective_level=eff, mode=mode, percentile=p,
                        numpy_semantics=sem, clamp_level=clamp, x=[float(v) for v in x.ravel()],
                        packed_hex=packed.tobytes().hex(), pruned_hex=pruned[0].numpy().tobytes().hex(), zeros=int(zeros)))
    path = os.path.join(ROOT, 'tests', 'golden', 'oracle_golden.json')
    with open(path, 'w') as f:
        json.dump(out, f)
    print(len(out), 'cases ->', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
