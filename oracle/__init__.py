"""oracle -- TEST INFRASTRUCTURE ONLY (CPU restatement of the reference's DWT-prune path).

PARITY UNPINNED: PyWavelets (the reference's dependency that holds the arithmetic) is neither
vendored in /root/reference nor installable here; see oracle/dwt_oracle.c for what pins it.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  wavelettransforms_b200 never does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

MODES = {'periodization': 0, 'symmetric': 1, 'zero': 2, 'constant': 3, 'reflect': 4, 'periodic': 5, 'smooth': 6,
         'antisymmetric': 7, 'antireflect': 8}


def build(force=False):
    so = os.path.join(_HERE, 'liboracle.so')
    src = os.path.join(_HERE, 'dwt_oracle.c')
    newest = max(os.path.getmtime(src), os.path.getmtime(os.path.join(_HERE, 'wavelet_tables.h')))
    if force or not os.path.exists(so) or os.path.getmtime(so) < newest:
        subprocess.check_call(['make', '-C', _HERE, '-s'])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, 'liboracle.so')
        if not os.path.exists(so):
            build()
        L = ctypes.CDLL(so)
        fp = ctypes.POINTER(ctypes.c_float)
        i64 = ctypes.c_int64
        ci = ctypes.c_int
        L.orc_wavelet_index.argtypes = [ctypes.c_char_p]
        L.orc_dec_len.argtypes = [ci]
        L.orc_dwt_max_level.argtypes = [ci, ci]
        L.orc_dwt_coeff_len.argtypes = [ci, ci, ci]
        L.orc_idwt_len.argtypes = [ci, ci, ci]
        L.orc_packed_side.argtypes = [ci, ci, ci, ci]
        L.orc_rec_size.argtypes = [ci, ci, ci, ci]
        L.orc_dwt.argtypes = [fp, ci, ci, ci, fp, fp]
        L.orc_idwt.argtypes = [fp, fp, ci, ci, ci, fp]
        L.orc_extend.argtypes = [fp, ci, ci, ci, ci, fp]
        L.orc_wavedec2_packed.argtypes = [fp, i64, ci, ci, ci, ci, ci, fp]
        L.orc_waverec2_packed.argtypes = [fp, i64, ci, ci, ci, ci, ci, fp]
        L.orc_percentile_plan.argtypes = [i64, ctypes.c_double, ci, ctypes.POINTER(i64),
                                          ctypes.POINTER(i64), ctypes.POINTER(ctypes.c_double)]
        L.orc_percentile_lerp.argtypes = [ctypes.c_float, ctypes.c_float, ctypes.c_double, ci]
        L.orc_percentile_lerp.restype = ctypes.c_float
        L.orc_percentile_threshold.argtypes = [fp, i64, ctypes.c_double, ci]
        L.orc_percentile_threshold.restype = ctypes.c_float
        L.orc_absolute_threshold.argtypes = [fp, i64, ctypes.c_float]
        L.orc_filters.argtypes = [ci] + [ctypes.POINTER(ctypes.c_double)] * 4
        L.orc_set_arithmetic.argtypes = [ci]
        _LIB = L
    return _LIB


ARITHMETICS = {'faithful': 0, 'fma_taporder': 1}


class arithmetic:
    """with oracle.arithmetic('fma_taporder'): ...  -- the oracle twin of the library's named fast mode (same tap
    order, every multiply-add fused); 'faithful' (no FMA) is the default and the reference's arithmetic"""

    def __init__(self, name):
        self.value = ARITHMETICS[name]

    def __enter__(self):
        self.prev = lib().orc_get_arithmetic()
        lib().orc_set_arithmetic(self.value)
        return self

    def __exit__(self, *exc):
        lib().orc_set_arithmetic(self.prev)
        return False


def _fp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def wavelet_index(name):
    wid = lib().orc_wavelet_index(str(name).encode())
    if wid < 0:
        raise ValueError(f"Unknown wavelet name '{name}', check wavelist() for the list of available builtin wavelets.")
    return wid


def mode_index(mode):
    if mode not in MODES:
        raise ValueError(f"Unknown mode name '{mode}'.")
    return MODES[mode]


def dec_len(wavelet):
    return lib().orc_dec_len(wavelet_index(wavelet))


def filters(wavelet):
    wid = wavelet_index(wavelet)
    bufs = [(ctypes.c_double * 18)() for _ in range(4)]
    n = lib().orc_filters(wid, *bufs)
    return tuple(np.array(b[:n]) for b in bufs)


def dwt_max_level(data_len, filter_len):
    return lib().orc_dwt_max_level(int(data_len), int(filter_len))


def dwt(x, wavelet, mode='symmetric'):
    x = np.ascontiguousarray(x, dtype=np.float32)
    wid, m = wavelet_index(wavelet), mode_index(mode)
    n = lib().orc_dwt_coeff_len(x.size, lib().orc_dec_len(wid), m)
    ca, cd = np.empty(n, np.float32), np.empty(n, np.float32)
    lib().orc_dwt(_fp(x), x.size, wid, m, _fp(ca), _fp(cd))
    return ca, cd


def idwt(ca, cd, wavelet, mode='symmetric'):
    ca = np.ascontiguousarray(ca, dtype=np.float32)
    cd = np.ascontiguousarray(cd, dtype=np.float32)
    wid, m = wavelet_index(wavelet), mode_index(mode)
    n = lib().orc_idwt_len(ca.size, lib().orc_dec_len(wid), m)
    out = np.empty(n, np.float32)
    r = lib().orc_idwt(_fp(ca), _fp(cd), ca.size, wid, m, _fp(out))
    if r < 0:
        raise ValueError('idwt: bad sizes')
    return out


def extend(x, before, after, mode):
    """pywt.pad(x, (before, after), mode): the boundary-extended signal of a non-periodization mode"""
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.size + before + after, np.float32)
    if lib().orc_extend(_fp(x), x.size, mode_index(mode), int(before), int(after), _fp(out)) < 0:
        raise ValueError('extend: bad arguments')
    return out


def packed_shape(shape, wavelet, level, mode='periodization'):
    wid, m = wavelet_index(wavelet), mode_index(mode)
    F = lib().orc_dec_len(wid)
    H, W = shape[-2:]
    return tuple(shape[:-2]) + (lib().orc_packed_side(H, F, m, level), lib().orc_packed_side(W, F, m, level))


def num_coeffs(shape, wavelet, level, mode='periodization'):
    """number of real coefficients per plane (the rest of the packed array is padding zeros)"""
    wid, m = wavelet_index(wavelet), mode_index(mode)
    F = lib().orc_dec_len(wid)
    H, W = shape[-2:]
    if level == 0:
        return H * W
    hs, ws = [H], [W]
    for _ in range(level):
        hs.append(lib().orc_dwt_coeff_len(hs[-1], F, m))
        ws.append(lib().orc_dwt_coeff_len(ws[-1], F, m))
    return hs[level] * ws[level] + sum(3 * hs[l] * ws[l] for l in range(1, level + 1))


def wavedec2_packed(x, wavelet, level, mode='periodization'):
    """pywt.wavedec2(x, wavelet, level=level, mode=mode, axes=(-2,-1)) followed by
    pywt.coeffs_to_array(coeffs, axes=(-2,-1))[0]   (dwt_pruning.py:67-70)."""
    if level < 0:
        raise ValueError(f"Level value of {level} is too low . Minimum level is 0.")
    x = np.ascontiguousarray(x, dtype=np.float32)
    wid, m = wavelet_index(wavelet), mode_index(mode)
    H, W = x.shape[-2:]
    batch = int(np.prod(x.shape[:-2], dtype=np.int64))
    out = np.empty(packed_shape(x.shape, wavelet, level, mode), np.float32)
    r = lib().orc_wavedec2_packed(_fp(x), batch, H, W, wid, level, m, _fp(out))
    if r:
        raise ValueError(f'orc_wavedec2_packed failed ({r})')
    return out


def waverec2_packed(packed, orig_shape, wavelet, level, mode='periodization'):
    """pywt.array_to_coeffs(...) + pywt.waverec2(...)   (dwt_pruning.py:75-77); uncropped."""
    packed = np.ascontiguousarray(packed, dtype=np.float32)
    wid, m = wavelet_index(wavelet), mode_index(mode)
    F = lib().orc_dec_len(wid)
    H, W = orig_shape[-2:]
    batch = int(np.prod(orig_shape[:-2], dtype=np.int64))
    RH, RW = lib().orc_rec_size(H, F, m, level), lib().orc_rec_size(W, F, m, level)
    out = np.empty(tuple(orig_shape[:-2]) + (RH, RW), np.float32)
    r = lib().orc_waverec2_packed(_fp(packed), batch, H, W, wid, level, m, _fp(out))
    if r:
        raise ValueError(f'orc_waverec2_packed failed ({r})')
    return out


def percentile_plan(n, p, semantics):
    k0, k1, g = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_double()
    lib().orc_percentile_plan(int(n), float(p), int(semantics), ctypes.byref(k0), ctypes.byref(k1), ctypes.byref(g))
    return k0.value, k1.value, g.value


def percentile_lerp(a, b, gamma, semantics):
    return np.float32(lib().orc_percentile_lerp(float(a), float(b), float(gamma), int(semantics)))
