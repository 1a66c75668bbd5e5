#!/usr/bin/env python
"""Derive the filter banks of the 17 wavelets named by the reference's --wavelet flag
(/root/reference/ResNet/main_pruning.py:89-96) and emit them as C tables.

PyWavelets (the reference's dependency, pinned 1.4.1 in requirements.txt:3) is not installed
here, so nothing is copied from it: orthogonal families are re-derived in 60-digit arithmetic
(Daubechies spectral factorisation; symlet root selection; coiflet moment conditions solved by
Gauss-Newton) and the biorthogonal spline/CDF filters come from their closed forms.  The
orientation conventions (which of the four filters is the reversed one, sign pattern of the
high-pass filters, zero padding of bior filters to a common even length) follow SURVEY.md
Appendix A.2.

Writes  oracle/wavelet_tables.h  and  wavelettransforms_b200/csrc/wt_wavelet_tables.h
(two independent copies so the product never includes anything under oracle/).
"""
import itertools
import os
import sys

import mpmath as mp

mp.mp.dps = 60
SQ2 = mp.sqrt(2)


def poly_mul(a, b):
    out = [mp.mpf(0)] * (len(a) + len(b) - 1)
    for i, x in enumerate(a):
        for j, y in enumerate(b):
            out[i + j] += x * y
    return out


def daubechies_root_groups(N):
    """Roots of the degree-(N-1) Daubechies polynomial mapped to the z plane.
    Returns a list of groups; each group is (roots_inside_unit_circle) for one real root or
    one conjugate pair, so a group can be flipped to its reciprocal as a whole."""
    if N == 1:
        return []
    coeffs = [mp.binomial(N - 1 + k, k) for k in range(N)]  # ascending in y
    ys = mp.polyroots(list(reversed(coeffs)), maxsteps=500, extraprec=400)
    groups, used = [], [False] * len(ys)
    for i, y in enumerate(ys):
        if used[i]:
            continue
        used[i] = True
        # y = (2 - z - 1/z)/4  ->  z^2 - (2-4y) z + 1 = 0
        b = 2 - 4 * y
        disc = mp.sqrt(b * b - 4)
        z1, z2 = (b + disc) / 2, (b - disc) / 2
        zin = z1 if abs(z1) < 1 else z2
        if abs(mp.im(y)) < mp.mpf(10) ** -40:
            groups.append([mp.mpc(mp.re(zin), 0)])
        else:
            # find the conjugate partner
            for j in range(i + 1, len(ys)):
                if not used[j] and abs(ys[j] - mp.conj(y)) < mp.mpf(10) ** -30:
                    used[j] = True
                    break
            groups.append([zin, mp.conj(zin)])
    return groups


def filter_from_roots(N, roots):
    """h(z) = c (1+z)^N prod (z - r), normalised to sum sqrt(2); returns real coeff list."""
    p = [mp.mpf(1)]
    for _ in range(N):
        p = poly_mul(p, [mp.mpf(1), mp.mpf(1)])
    p = [mp.mpc(x) for x in p]
    for r in roots:
        p = poly_mul(p, [-r, mp.mpc(1)])
    p = [mp.re(x) for x in p]
    s = sum(p)
    return [x * SQ2 / s for x in p]


def daubechies(N):
    """rec_lo of dbN: minimum-phase choice, largest coefficients first."""
    roots = [r for g in daubechies_root_groups(N) for r in g]
    h = filter_from_roots(N, roots)
    # min-phase polynomial in ascending powers has its energy at the front already
    if sum(abs(x) for x in h[: N]) < sum(abs(x) for x in h[N:]):
        h = h[::-1]
    return h


def symlet(N, approx_rec_lo):
    """rec_lo of symN: the root selection whose filter is closest to the well-known
    least-asymmetric table values (approx_rec_lo, ~8 digits are enough to pick it)."""
    groups = daubechies_root_groups(N)
    best, best_err = None, None
    for flips in itertools.product([0, 1], repeat=len(groups)):
        roots = []
        for f, g in zip(flips, groups):
            roots += [1 / r if f else r for r in g]
        h = filter_from_roots(N, roots)
        for cand in (h, h[::-1]):
            err = max(abs(a - mp.mpf(b)) for a, b in zip(cand, approx_rec_lo))
            if best_err is None or err < best_err:
                best, best_err = cand, err
    assert best_err < mp.mpf('1e-4'), best_err
    return best


def coiflet(N, approx_dec_lo):
    """dec_lo of coifN (length 6N) by Gauss-Newton on orthonormality + 2N wavelet vanishing
    moments + (2N-1) scaling-function vanishing moments, started from table values."""
    F = 6 * N
    h = mp.matrix([mp.mpf(x) for x in approx_dec_lo])
    c = int(mp.nint(sum(n * h[n] for n in range(F)) / SQ2))

    def residual(h):
        res = []
        for k in range(F // 2):
            s = sum(h[n] * h[n + 2 * k] for n in range(F - 2 * k))
            res.append(s - (1 if k == 0 else 0))
        res.append(sum(h[n] for n in range(F)) - SQ2)
        for m in range(2 * N):
            res.append(sum((-1) ** n * mp.mpf(n - c) ** m * h[n] for n in range(F)))
        for m in range(1, 2 * N):
            res.append(sum(mp.mpf(n - c) ** m * h[n] for n in range(F)))
        return mp.matrix(res)

    def jac(h):
        rows = []
        for k in range(F // 2):
            row = [mp.mpf(0)] * F
            for n in range(F - 2 * k):
                row[n] += h[n + 2 * k]
                row[n + 2 * k] += h[n]
            rows.append(row)
        rows.append([mp.mpf(1)] * F)
        for m in range(2 * N):
            rows.append([(-1) ** n * mp.mpf(n - c) ** m for n in range(F)])
        for m in range(1, 2 * N):
            rows.append([mp.mpf(n - c) ** m for n in range(F)])
        return mp.matrix(rows)

    for _ in range(60):
        r = residual(h)
        J = jac(h)
        step = mp.lu_solve(J.T * J, -(J.T * r))
        h = h + step
        if mp.norm(step) < mp.mpf(10) ** -50:
            break
    assert mp.norm(residual(h)) < mp.mpf(10) ** -40, mp.norm(residual(h))
    err = max(abs(h[n] - mp.mpf(approx_dec_lo[n])) for n in range(F))
    assert err < mp.mpf('1e-6'), err
    return [h[n] for n in range(F)]


def orth_bank(rec_lo):
    F = len(rec_lo)
    dec_lo = rec_lo[::-1]
    rec_hi = [(-1) ** i * rec_lo[F - 1 - i] for i in range(F)]
    dec_hi = rec_hi[::-1]
    return dec_lo, dec_hi, rec_lo, rec_hi


def bior_banks():
    a, b = SQ2 / 16, SQ2 / 2
    c, d, e = SQ2 / 8, SQ2 / 4, 3 * SQ2 / 4
    z = mp.mpf(0)
    banks = {
        'bior1.3': ([-a, a, b, b, a, -a], [z, z, -b, b, z, z], [z, z, b, b, z, z], [-a, -a, b, -b, a, a]),
        'bior2.2': ([z, -c, d, e, d, -c], [z, d, -b, d, z, z], [z, d, b, d, z, z], [z, c, d, -e, d, c]),
    }
    # CDF 9/7 (bior4.4).  Closed form through the real root of the degree-3 Daubechies
    # polynomial: the 9-tap and 7-tap symmetric factors of the N=4 half-band product.
    groups = daubechies_root_groups(4)
    real = [g for g in groups if len(g) == 1][0][0]
    cplx = [g for g in groups if len(g) == 2][0]
    # 7-tap synthesis low-pass: (1+z)^4 * (z-r)(z-1/r)   (real root pair)
    p7 = [mp.mpf(1)]
    for _ in range(4):
        p7 = poly_mul(p7, [mp.mpf(1), mp.mpf(1)])
    p7 = poly_mul(p7, [mp.re(real) * 0 + 1, -(mp.re(real) + 1 / mp.re(real)), mp.mpf(1)])
    s = sum(p7)
    p7 = [x * SQ2 / s for x in p7]
    # 9-tap analysis low-pass: (1+z)^4 * (z-c)(z-c*)(z-1/c)(z-1/c*)
    p9 = [mp.mpc(1)]
    for _ in range(4):
        p9 = poly_mul(p9, [mp.mpc(1), mp.mpc(1)])
    for r in (cplx[0], cplx[1], 1 / cplx[0], 1 / cplx[1]):
        p9 = poly_mul(p9, [-r, mp.mpc(1)])
    p9 = [mp.re(x) for x in p9]
    s = sum(p9)
    p9 = [x * SQ2 / s for x in p9]
    dec_lo = [z] + p9                       # length 10
    rec_lo = [z] + p7 + [z, z]              # length 10
    # high-pass by alternating-sign modulation of the opposite low-pass
    # rec_hi[i] = (-1)^i dec_lo[i] ; dec_hi[i] = (-1)^(i+1) rec_lo[i]  (same rule bior1.3/2.2 obey)
    rec_hi = [(-1) ** i * dec_lo[i] for i in range(10)]
    dec_hi = [(-1) ** (i + 1) * rec_lo[i] for i in range(10)]
    banks['bior4.4'] = (dec_lo, dec_hi, rec_lo, rec_hi)
    for name in list(banks):
        dl, dh, rl, rh = banks[name]
        banks[name.replace('bior', 'rbio')] = (rl[::-1], rh[::-1], dl[::-1], dh[::-1])
    return banks


SYM4_REC_LO = [0.0322231006040427, -0.012603967262037833, -0.09921954357684722, 0.29785779560527736,
               0.8037387518059161, 0.49761866763201545, -0.02963552764599851, -0.07576571478927333]
SYM6_DEC_LO = [0.015404109327027373, 0.0034907120842174702, -0.11799011114819057, -0.048311742585633,
               0.4910559419267466, 0.787641141030194, 0.3379294217276218, -0.07263752278646252,
               -0.021060292512300564, 0.0447249017706658, 0.0017677118642428036, -0.007800708325034148]
COIF1_DEC_LO = [-0.01565572813546454, -0.0727326195128539, 0.38486484686420286, 0.8525720202122554,
                0.3378976624578092, -0.0727326195128539]
COIF2_DEC_LO = [-0.0007205494453645122, -0.0018232088707029932, 0.0056114348193944995, 0.023680171946334084,
                -0.0594344186464569, -0.0764885990783064, 0.41700518442169254, 0.8127236354455423,
                0.3861100668211622, -0.06737255472196302, -0.04146493678175915, 0.016387336463522112]
COIF3_DEC_LO = [-3.459977283621256e-05, -7.098330313814125e-05, 0.0004662169601128863, 0.0011175187708906016,
                -0.0025745176887502236, -0.00900797613666158, 0.015880544863615904, 0.03455502757306163,
                -0.08230192710688598, -0.07179982161931202, 0.42848347637761874, 0.7937772226256206,
                0.4051769024096169, -0.06112339000267287, -0.0657719112818555, 0.023452696141836267,
                0.007782596427325418, -0.003793512864491014]


def all_banks():
    banks = {}
    banks['haar'] = orth_bank(daubechies(1))
    banks['db1'] = banks['haar']
    for n in (2, 4, 6):
        banks[f'db{n}'] = orth_bank(daubechies(n))
    banks['sym2'] = banks['db2']
    banks['sym4'] = orth_bank(symlet(4, SYM4_REC_LO))
    banks['sym6'] = orth_bank(symlet(6, SYM6_DEC_LO[::-1]))
    banks['coif1'] = orth_bank(coiflet(1, COIF1_DEC_LO)[::-1])
    banks['coif2'] = orth_bank(coiflet(2, COIF2_DEC_LO)[::-1])
    banks['coif3'] = orth_bank(coiflet(3, COIF3_DEC_LO)[::-1])
    banks.update(bior_banks())
    # round 2 (`wavelet` is an open string in dwt_pruning.py:35): the Daubechies filters up to 18 taps need
    # nothing beyond the spectral factorisation above; sym3 is db3 (below N = 4 the least-asymmetric choice is the only one);
    # bior1.1 is haar; bior3.1 is the 4/4-tap spline pair in closed form.  The other symlets / coiflets / spline pairs need
    # a table to pick their root selection or alignment from and stay out until PyWavelets can be held against them.
    for n in (3, 5, 7, 8, 9):
        banks[f'db{n}'] = orth_bank(daubechies(n))
    banks['sym3'] = banks['db3']
    banks['bior1.1'] = banks['haar']
    banks['rbio1.1'] = tuple(f[::-1] for f in (banks['haar'][2], banks['haar'][3], banks['haar'][0], banks['haar'][1]))
    q4, q8 = SQ2 / 4, SQ2 / 8
    b31_dec_lo = [-q4, 3 * q4, 3 * q4, -q4]
    b31_rec_lo = [q8, 3 * q8, 3 * q8, q8]
    b31 = (b31_dec_lo, [(-1) ** (i + 1) * b31_rec_lo[i] for i in range(4)], b31_rec_lo, [(-1) ** i * b31_dec_lo[i] for i in range(4)])
    banks['bior3.1'] = b31
    banks['rbio3.1'] = (b31[2][::-1], b31[3][::-1], b31[0][::-1], b31[1][::-1])
    return banks


ORDER = ['haar', 'db1', 'db2', 'db4', 'db6', 'coif1', 'coif2', 'coif3', 'bior1.3', 'bior2.2', 'bior4.4',
         'rbio1.3', 'rbio2.2', 'rbio4.4', 'sym2', 'sym4', 'sym6',
         # round 2, appended so that the ids of the 17 flag names stay what they were
         'db3', 'db5', 'db7', 'db8', 'db9', 'sym3', 'bior1.1', 'rbio1.1', 'bior3.1', 'rbio3.1']
MAXF = 18


def fmt(x):
    x = float(x)
    return repr(x if x != 0 else 0.0)


def emit(banks, guard, prefix):
    lines = [f'/* GENERATED by tools/gen_wavelets.py -- do not edit.  Filter banks of the reference\'s',
             ' * --wavelet enum (/root/reference/ResNet/main_pruning.py:89-96); conventions per SURVEY.md A.2. */',
             f'#ifndef {guard}', f'#define {guard}',
             f'#define {prefix}_NUM_WAVELETS {len(ORDER)}', f'#define {prefix}_MAX_FILTER_LEN {MAXF}',
             f'typedef struct {{ const char* name; int len; double dec_lo[{MAXF}], dec_hi[{MAXF}], rec_lo[{MAXF}], rec_hi[{MAXF}]; }} {prefix.lower()}_bank_t;',
             f'static const {prefix.lower()}_bank_t {prefix}_BANKS[{len(ORDER)}] = {{']
    for name in ORDER:
        dl, dh, rl, rh = banks[name]
        body = ',\n    '.join('{' + ', '.join(fmt(v) for v in f) + '}' for f in (dl, dh, rl, rh))
        lines.append(f'  {{"{name}", {len(dl)},\n    {body}}},')
    lines += ['};', f'#endif /* {guard} */', '']
    return '\n'.join(lines)


def check_pr(banks):
    """Perfect reconstruction of the two-channel bank: sum_k dec*rec products."""
    import numpy as np
    worst = 0.0
    for name in ORDER:
        dl, dh, rl, rh = [np.array([float(v) for v in f]) for f in banks[name]]
        # polyphase PR test: rec_lo*dec_lo + rec_hi*dec_hi must be 2*delta at odd centre, and the
        # aliasing term must vanish.
        t = np.convolve(rl, dl) + np.convolve(rh, dh)
        alt = (-1.0) ** np.arange(len(dl))
        al = np.convolve(rl, dl * alt) + np.convolve(rh, dh * alt)
        centre = len(dl) - 1
        want = np.zeros_like(t)
        want[centre] = 2.0
        worst = max(worst, np.abs(t - want).max(), np.abs(al).max())
    return worst


def main():
    banks = all_banks()
    w = check_pr(banks)
    print('worst PR residual', w)
    assert w < 1e-14
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(os.path.join(root, 'oracle', 'wavelet_tables.h'), 'w') as f:
        f.write(emit(banks, 'ORACLE_WAVELET_TABLES_H', 'ORC'))
    with open(os.path.join(root, 'wavelettransforms_b200', 'csrc', 'wt_wavelet_tables.h'), 'w') as f:
        f.write(emit(banks, 'WT_WAVELET_TABLES_H', 'WT'))
    for name in ORDER:
        print(name, len(banks[name][0]), [float(v) for v in banks[name][0]][:4])


if __name__ == '__main__':
    sys.exit(main())
