"""ctypes binding of libwtprune.so (include/wtprune.h).  No CPU fallback: a missing library or a
missing CUDA device raises."""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, 'lib', 'libwtprune.so')

WT_OK, WT_E_INVALID, WT_E_CUDA, WT_E_WORKSPACE, WT_E_UNSUPPORTED = 0, -1, -2, -3, -4
# pywt.Modes.modes; ids are include/wtprune.h's WT_MODE_*
MODES = {'periodization': 0, 'symmetric': 1, 'zero': 2, 'constant': 3, 'reflect': 4, 'periodic': 5, 'smooth': 6,
         'antisymmetric': 7, 'antireflect': 8}
NUMPY_SEMANTICS = {'legacy_1_22': 0, 'nep50_2x': 1}
THRESHOLD_PERCENTILE, THRESHOLD_ABSOLUTE = 0, 1
ARITHMETICS = {'faithful': 0, 'fma_taporder': 1}


class WtLayerDesc(ctypes.Structure):
    _fields_ = [('in_', ctypes.c_void_p), ('out', ctypes.c_void_p), ('batch', ctypes.c_int64),
                ('H', ctypes.c_int32), ('W', ctypes.c_int32), ('level', ctypes.c_int32), ('reserved', ctypes.c_int32)]


# every symbol include/wtprune.h declares (tests check the library exports each of them)
EXPORTS = [
    'wt_version', 'wt_last_error', 'wt_wavelet_lookup', 'wt_wavelet_filters', 'wt_dwt_max_level', 'wt_dwt_coeff_len',
    'wt_packed_shape', 'wt_percentile_plan', 'wt_plan_create', 'wt_plan_destroy', 'wt_plan_device_bytes',
    'wt_plan_num_launches', 'wt_plan_set_io', 'wt_plan_run', 'wt_plan_results', 'wt_plan_counts_device', 'wt_plan_thresholds_device',
    'wt_plan_set_timing', 'wt_plan_phase_times', 'wt_plan_set_peers', 'wt_plan_select_path', 'wt_plan_set_arithmetic', 'wt_plan_set_multicast', 'wt_plan_set_graph', 'wt_plan_run_stage', 'wt_plan_layer_coefficients',
    'wt_dwt_prune_f32', 'wt_select_abs_f32', 'wt_select_abs_noalloc_f32', 'wt_wavedec2_packed_f32', 'wt_waverec2_packed_f32',
    'wt_min_weight_prune_f32', 'wt_scatter_zero_f32', 'wt_count_f32', 'wt_key_hist_f32',
]

_LIB = None


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f'{LIB_PATH} is missing: build it with `python -m wavelettransforms_b200.build` '
            '(there is no CPU fallback for this path)')
    L = ctypes.CDLL(LIB_PATH)
    vp, ci, i64, dbl, fl = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double, ctypes.c_float
    P = ctypes.POINTER
    L.wt_version.restype = ci
    L.wt_last_error.restype = ctypes.c_char_p
    L.wt_wavelet_lookup.argtypes = [ctypes.c_char_p, P(ci), P(ci)]
    L.wt_wavelet_filters.argtypes = [ci, P(dbl), P(dbl), P(dbl), P(dbl)]
    L.wt_dwt_max_level.argtypes = [ci, ci]
    L.wt_dwt_coeff_len.argtypes = [ci, ci, ci]
    L.wt_packed_shape.argtypes = [ci, ci, ci, ci, ci, P(ci), P(ci), P(i64)]
    L.wt_percentile_plan.argtypes = [i64, dbl, ci, P(i64), P(i64), P(dbl)]
    L.wt_plan_create.argtypes = [P(WtLayerDesc), ci, ci, ci, ci, dbl, ci, ci, P(vp)]
    L.wt_plan_destroy.argtypes = [vp]
    L.wt_plan_device_bytes.argtypes = [vp]
    L.wt_plan_device_bytes.restype = ctypes.c_size_t
    L.wt_plan_num_launches.argtypes = [vp]
    L.wt_plan_set_io.argtypes = [vp, P(WtLayerDesc), ci]
    L.wt_plan_run.argtypes = [vp, vp]
    L.wt_plan_results.argtypes = [vp, vp, P(i64), P(i64), P(fl)]
    L.wt_plan_counts_device.argtypes = [vp, P(vp)]
    L.wt_plan_thresholds_device.argtypes = [vp, P(vp)]
    L.wt_plan_set_timing.argtypes = [vp, ci]
    L.wt_plan_set_graph.argtypes = [vp, ci]
    L.wt_plan_run_stage.argtypes = [vp, ci, vp]
    L.wt_plan_layer_coefficients.argtypes = [vp, ci, P(vp), P(i64), P(i64)]
    L.wt_plan_select_path.argtypes = [vp]
    L.wt_plan_set_arithmetic.argtypes = [vp, ci]
    L.wt_plan_set_multicast.argtypes = [vp, i64]
    L.wt_plan_set_peers.argtypes = [vp, ci, ctypes.POINTER(ctypes.c_int64)]
    L.wt_plan_phase_times.argtypes = [vp, ci, P(fl), P(ci), ctypes.c_char_p, ci]
    L.wt_dwt_prune_f32.argtypes = [P(WtLayerDesc), ci, ci, ci, ci, dbl, ci, vp, P(i64), P(i64), P(fl)]
    L.wt_select_abs_f32.argtypes = [vp, i64, i64, i64, i64, vp, vp]
    L.wt_select_abs_noalloc_f32.argtypes = [vp, i64, i64, i64, i64, vp, vp]
    L.wt_wavedec2_packed_f32.argtypes = [vp, i64, ci, ci, ci, ci, ci, vp, vp]
    L.wt_waverec2_packed_f32.argtypes = [vp, i64, ci, ci, ci, ci, ci, vp, vp]
    L.wt_min_weight_prune_f32.argtypes = [P(vp), P(i64), P(i64), ci, vp, P(i64)]
    L.wt_scatter_zero_f32.argtypes = [vp, i64, vp, i64, vp, vp]
    L.wt_count_f32.argtypes = [vp, i64, fl, vp, vp]
    L.wt_key_hist_f32.argtypes = [vp, i64, ctypes.c_uint32, ci, ci, vp, vp]
    _LIB = L
    return L


def last_error():
    return lib().wt_last_error().decode(errors='replace')


def check(rc):
    """maps C status codes onto the exceptions the reference's dependencies raise: ValueError for
    bad wavelet / mode / level / percentile (pywt, numpy), RuntimeError for everything else."""
    if rc == WT_OK:
        return
    msg = last_error()
    if rc == WT_E_INVALID:
        raise ValueError(msg)
    if rc == WT_E_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise RuntimeError(f'libwtprune: {msg} (code {rc})')


def wavelet_lookup(name):
    wid, F = ctypes.c_int(), ctypes.c_int()
    check(lib().wt_wavelet_lookup(str(name).encode(), ctypes.byref(wid), ctypes.byref(F)))
    return wid.value, F.value


def wavelet_filters(name):
    wid, F = wavelet_lookup(name)
    bufs = [(ctypes.c_double * 18)() for _ in range(4)]
    check(lib().wt_wavelet_filters(wid, *bufs))
    return tuple(np.array(b[:F]) for b in bufs)


def mode_index(mode):
    if mode not in MODES:
        raise ValueError(f"Unknown mode name '{mode}'.")
    return MODES[mode]


def semantics_index(sem):
    if sem not in NUMPY_SEMANTICS:
        raise ValueError(f"unknown numpy_semantics '{sem}' (expected one of {sorted(NUMPY_SEMANTICS)})")
    return NUMPY_SEMANTICS[sem]


def dwt_max_level(data_len, filter_len):
    return lib().wt_dwt_max_level(int(data_len), int(filter_len))


def packed_shape(H, W, wavelet, level, mode='periodization'):
    wid, _ = wavelet_lookup(wavelet)
    sh, sw, nc = ctypes.c_int(), ctypes.c_int(), ctypes.c_int64()
    check(lib().wt_packed_shape(int(H), int(W), wid, int(level), mode_index(mode), ctypes.byref(sh), ctypes.byref(sw),
                                ctypes.byref(nc)))
    return sh.value, sw.value, nc.value


def percentile_plan(n, p, semantics):
    k0, k1, g = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_double()
    check(lib().wt_percentile_plan(int(n), float(p), semantics_index(semantics), ctypes.byref(k0), ctypes.byref(k1),
                                   ctypes.byref(g)))
    return k0.value, k1.value, g.value


def make_descs(items):
    """items: iterable of (in_ptr, out_ptr, batch, H, W, level)"""
    items = list(items)
    arr = (WtLayerDesc * max(len(items), 1))()
    for i, (pi, po, b, h, w, lv) in enumerate(items):
        arr[i].in_, arr[i].out, arr[i].batch, arr[i].H, arr[i].W, arr[i].level = pi, po, int(b), int(h), int(w), int(lv)
    return arr, len(items)


class Plan:
    """Owns a WtPlan*."""

    def __init__(self, items, wavelet, mode='periodization', threshold_kind=THRESHOLD_PERCENTILE, threshold_value=50.0,
                 numpy_semantics='legacy_1_22', host_io=False):
        wid, _ = wavelet_lookup(wavelet)
        self.n = len(list(items)) if not isinstance(items, list) else len(items)
        descs, n = make_descs(items)
        self.n = n
        self._h = ctypes.c_void_p()
        check(lib().wt_plan_create(descs, n, wid, mode_index(mode), int(threshold_kind), float(threshold_value),
                                   semantics_index(numpy_semantics), 1 if host_io else 0, ctypes.byref(self._h)))

    def set_io(self, items):
        descs, n = make_descs(items)
        check(lib().wt_plan_set_io(self._h, descs, n))

    def run(self, stream=0):
        check(lib().wt_plan_run(self._h, ctypes.c_void_p(stream)))

    def results(self, stream=0):
        n = max(self.n, 1)
        counts = (ctypes.c_int64 * (3 * n))()
        nzb = (ctypes.c_int64 * n)()
        thr = (ctypes.c_float * n)()
        check(lib().wt_plan_results(self._h, ctypes.c_void_p(stream), counts, nzb, thr))
        c = np.array(counts[:3 * self.n], dtype=np.int64).reshape(self.n, 3)
        return c, np.array(nzb[:self.n], dtype=np.int64), np.array(thr[:self.n], dtype=np.float32)

    def run_stage(self, stage, stream=0):
        """1: analysis only; 2: threshold (whatever the plan's threshold array holds) + synthesis (wtprune.h: wt_plan_run_stage)"""
        check(lib().wt_plan_run_stage(self._h, int(stage), ctypes.c_void_p(stream)))

    def layer_coefficients(self, layer):
        """(device address, explicit coefficients, implicit padding zeros) of a layer after run_stage(1)"""
        p, n, z = ctypes.c_void_p(), ctypes.c_int64(), ctypes.c_int64()
        check(lib().wt_plan_layer_coefficients(self._h, int(layer), ctypes.byref(p), ctypes.byref(n), ctypes.byref(z)))
        return int(p.value or 0), int(n.value), int(z.value)

    def device_result_pointers(self):
        """(address of the int64 [n][4] block {scratch, nonzero, zeros, nonzero_before} of the last run, address of the
        float32 [n] thresholds); valid in stream order after run(), nothing is synchronised"""
        c, t = ctypes.c_void_p(), ctypes.c_void_p()
        check(lib().wt_plan_counts_device(self._h, ctypes.byref(c)))
        check(lib().wt_plan_thresholds_device(self._h, ctypes.byref(t)))
        return int(c.value or 0), int(t.value or 0)

    def set_peers(self, byte_deltas):
        """mirror every output store at out + delta for each delta (peer mappings of the same flat buffer layout)"""
        d = [int(x) for x in byte_deltas]
        arr = (ctypes.c_int64 * max(len(d), 1))(*d)
        check(lib().wt_plan_set_peers(self._h, len(d), arr))

    def set_multicast(self, byte_delta):
        """mirror every output store once through the multicast (NVLS) mapping at out + byte_delta"""
        check(lib().wt_plan_set_multicast(self._h, int(byte_delta)))

    def set_arithmetic(self, name):
        """'faithful' (default: no FMA, the oracle's arithmetic) or 'fma_taporder' (named fast mode, see wtprune.h)"""
        if name not in ARITHMETICS:
            raise ValueError(f"unknown arithmetic '{name}' (expected one of {sorted(ARITHMETICS)})")
        check(lib().wt_plan_set_arithmetic(self._h, ARITHMETICS[name]))

    def set_graph(self, enable=True):
        """replay the plan's launch sequence as a CUDA graph from the third run on (wtprune.h: wt_plan_set_graph)"""
        check(lib().wt_plan_set_graph(self._h, 1 if enable else 0))

    def set_timing(self, enable=True):
        check(lib().wt_plan_set_timing(self._h, 1 if enable else 0))

    def phase_times(self, max_phases=256):
        """[(name, ms, kernel launches)] of the last run (waits for it); needs set_timing(True)"""
        ms = (ctypes.c_float * max_phases)()
        ln = (ctypes.c_int * max_phases)()
        names = ctypes.create_string_buffer(max_phases * 64)
        n = lib().wt_plan_phase_times(self._h, max_phases, ms, ln, names, 64)
        if n < 0:
            check(n)
        raw = names.raw
        return [(raw[i * 64:(i + 1) * 64].split(b'\0', 1)[0].decode(), float(ms[i]), int(ln[i])) for i in range(n)]

    @property
    def select_path(self):
        return {0: 'sampled', 1: 'radix', 2: 'fused', 3: 'precompacted', -1: None}[lib().wt_plan_select_path(self._h)]

    @property
    def device_bytes(self):
        return lib().wt_plan_device_bytes(self._h)

    @property
    def num_launches(self):
        return lib().wt_plan_num_launches(self._h)

    def close(self):
        if self._h:
            lib().wt_plan_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
