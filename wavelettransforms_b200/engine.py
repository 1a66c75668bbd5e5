"""Host-side engine: torch tensors in, libwtprune plans out.  PyTorch is used for device memory,
streams and the host<->device copies only; every numeric step runs in libwtprune's kernels."""
import ctypes
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib, config


def cuda_device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError('wavelettransforms_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback')
    return torch.device('cuda', torch.cuda.current_device())


def _stream_ptr() -> int:
    return int(torch.cuda.current_stream().cuda_stream)


def plane_geometry(shape: Sequence[int]) -> Tuple[int, int, int]:
    """(batch, H, W) of a tensor transformed over its last two axes; <2-D tensors are one row."""
    shape = tuple(int(s) for s in shape)
    if len(shape) < 2:
        n = int(np.prod(shape, dtype=np.int64)) if len(shape) else 1
        return 1, 1, n
    return int(np.prod(shape[:-2], dtype=np.int64)), shape[-2], shape[-1]


def calculate_max_level(shape, wavelet) -> int:
    """pywt.dwt_max_level(min(shape[-2:]), pywt.Wavelet(wavelet).dec_len)  (dwt_pruning.py:12-13)"""
    _, dec_len = _lib.wavelet_lookup(wavelet)
    return _lib.dwt_max_level(min(shape[-2:]), dec_len)


def _check_fp32(t: torch.Tensor):
    if t.dtype != torch.float32:
        raise NotImplementedError(
            f'wavelettransforms_b200 computes in float32 like the reference does for float32 weights; got {t.dtype}')


def to_device_f32(t: torch.Tensor, dev: torch.device) -> torch.Tensor:
    _check_fp32(t)
    return t.detach().to(dev, non_blocking=False).contiguous()


class PruneResult:
    __slots__ = ('outputs', 'counts', 'nonzero_before', 'thresholds', 'launches')

    def __init__(self, outputs, counts, nonzero_before, thresholds, launches):
        self.outputs = outputs                # list of device tensors, same shapes as the inputs
        self.counts = counts                  # int64 [n, 3]: numel, nonzero after, zeros after
        self.nonzero_before = nonzero_before  # int64 [n]
        self.thresholds = thresholds          # float32 [n]
        self.launches = launches


class DevicePruneResult:
    """prune_device_tensors(..., device_results=True): the same numbers as PruneResult, left on the GPU.  No D2H copy
    has been made and the tensors are valid in stream order; .host() does the one D2H + sync when the caller wants it.
    (Tearing the one-shot plan down still waits for the device inside cudaFree; a caller that must not block at all
    keeps a _lib.Plan alive and reads Plan.device_result_pointers() after each run.)"""
    __slots__ = ('outputs', 'numel', 'counts_dev', 'thresholds_dev', 'launches')

    def __init__(self, outputs, numel, counts_dev, thresholds_dev, launches):
        self.outputs = outputs                # list of device tensors
        self.numel = numel                    # int64 [n] (host: known without running anything)
        self.counts_dev = counts_dev          # int64 CUDA tensor [n, 3]: nonzero after, zeros after, nonzero before
        self.thresholds_dev = thresholds_dev  # float32 CUDA tensor [n]
        self.launches = launches

    def host(self) -> PruneResult:
        c = self.counts_dev.cpu().numpy()
        counts = np.stack([self.numel, c[:, 0], c[:, 1]], axis=1) if len(self.numel) else np.zeros((0, 3), dtype=np.int64)
        return PruneResult(self.outputs, counts, c[:, 2].copy(), self.thresholds_dev.cpu().numpy(), self.launches)


class _DeviceBlock:
    """a raw device address as a __cuda_array_interface__ object, so torch can view it without a copy"""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {'shape': tuple(shape), 'typestr': typestr, 'data': (int(ptr), False), 'version': 2}


def prune_device_tensors(xs: List[torch.Tensor], levels: Sequence[int], wavelet: str, mode: str, threshold_kind: int,
                         threshold_value: float, numpy_semantics: Optional[str] = None,
                         outs: Optional[List[torch.Tensor]] = None,
                         peer_deltas: Optional[Sequence[int]] = None, arithmetic: Optional[str] = None,
                         device_results: bool = False):
    """xs: contiguous float32 CUDA tensors.  One plan, one launch sequence for the whole group.  peer_deltas: the
    kernels also store every output at out + delta (symmetric-memory mappings of the same buffer on the other ranks).
    device_results: return a DevicePruneResult (counts and thresholds stay on the GPU, no stream synchronisation)."""
    sem = numpy_semantics or config.numpy_semantics
    for lv in levels:
        if lv < 0:
            raise ValueError(f'Level value of {lv} is too low . Minimum level is 0.')
    if outs is None:
        outs = [torch.empty_like(x) for x in xs]
    items = []
    for x, o, lv in zip(xs, outs, levels):
        b, h, w = plane_geometry(x.shape)
        if x.ndim < 2:
            lv = 0
        items.append((x.data_ptr(), o.data_ptr(), b, h, w, lv))
    plan = _lib.Plan(items, wavelet, mode, threshold_kind, threshold_value, sem, host_io=False)
    try:
        arith = arithmetic or config.arithmetic
        if arith != 'faithful':
            plan.set_arithmetic(arith)
        if isinstance(peer_deltas, tuple) and peer_deltas and peer_deltas[0] == 'multicast':
            plan.set_multicast(peer_deltas[1])
        elif peer_deltas:
            plan.set_peers(peer_deltas)
        s = _stream_ptr()
        plan.run(s)
        launches = plan.num_launches
        if device_results:
            n = len(items)
            numel = np.array([b * h * w for _, _, b, h, w, _ in items], dtype=np.int64)
            if n == 0:
                dev = cuda_device()
                return DevicePruneResult(outs, numel, torch.zeros((0, 3), dtype=torch.int64, device=dev),
                                         torch.zeros(0, dtype=torch.float32, device=dev), launches)
            cp, tp = plan.device_result_pointers()
            dev = xs[0].device
            # copies on the current stream, enqueued before plan.close() frees the plan's block (cudaFree waits for them)
            c4 = torch.as_tensor(_DeviceBlock(cp, (n, 4), '<i8'), device=dev)
            counts_dev = c4[:, 1:].clone()
            thresholds_dev = torch.as_tensor(_DeviceBlock(tp, (n,), '<f4'), device=dev).clone()
            return DevicePruneResult(outs, numel, counts_dev, thresholds_dev, launches)
        counts, nzb, thr = plan.results(s)
    finally:
        plan.close()
    return PruneResult(outs, counts, nzb, thr, launches)


# ---- building blocks -------------------------------------------------------------------------------

def select_abs(x: torch.Tensor, k0: int, k1: int, implicit_zeros: int = 0) -> Tuple[float, float]:
    """sorted(|x| U zeros)[k0], [k1] -- the partition step of np.percentile"""
    x = x.contiguous()
    _check_fp32(x)
    ab = torch.empty(2, dtype=torch.float32, device=x.device)
    _lib.check(_lib.lib().wt_select_abs_f32(x.data_ptr(), x.numel(), int(implicit_zeros), int(k0), int(k1), ab.data_ptr(),
                                            _stream_ptr()))
    a, b = ab.cpu().tolist()
    return a, b


def wavedec2_packed(x: torch.Tensor, wavelet: str, level: int, mode: str = 'periodization') -> torch.Tensor:
    """pywt.coeffs_to_array(pywt.wavedec2(x, wavelet, level=level, mode=mode, axes=(-2,-1)))[0] on the GPU"""
    _check_fp32(x)
    x = x.contiguous()
    wid, _ = _lib.wavelet_lookup(wavelet)
    if level < 0:
        raise ValueError(f'Level value of {level} is too low . Minimum level is 0.')
    b, h, w = plane_geometry(x.shape)
    sh, sw, _ = _lib.packed_shape(h, w, wavelet, level, mode)
    out = torch.empty(tuple(x.shape[:-2]) + (sh, sw), dtype=torch.float32, device=x.device)
    _lib.check(_lib.lib().wt_wavedec2_packed_f32(x.data_ptr(), b, h, w, wid, int(level), _lib.mode_index(mode), out.data_ptr(),
                                                 _stream_ptr()))
    return out


def waverec2_packed(packed: torch.Tensor, orig_shape, wavelet: str, level: int, mode: str = 'periodization') -> torch.Tensor:
    """pywt.waverec2(pywt.array_to_coeffs(packed, slices, 'wavedec2'), wavelet, mode) cropped to orig_shape"""
    _check_fp32(packed)
    packed = packed.contiguous()
    wid, _ = _lib.wavelet_lookup(wavelet)
    b, h, w = plane_geometry(orig_shape)
    out = torch.empty(tuple(orig_shape), dtype=torch.float32, device=packed.device)
    _lib.check(_lib.lib().wt_waverec2_packed_f32(packed.data_ptr(), b, h, w, wid, int(level), _lib.mode_index(mode),
                                                 out.data_ptr(), _stream_ptr()))
    return out


# ---- sibling maskers --------------------------------------------------------------------------------

def min_weight_prune_(tensors: List[torch.Tensor], ks: Sequence[int]) -> List[int]:
    """in place on contiguous float32 CUDA tensors: zero the ks[i] smallest |w|; returns nonzero counts"""
    n = len(tensors)
    if n == 0:
        return []
    for t in tensors:
        _check_fp32(t)
        if not t.is_contiguous():
            raise ValueError('min_weight_prune_ needs contiguous tensors')
    ptrs = (ctypes.c_void_p * n)(*[t.data_ptr() for t in tensors])
    ns = (ctypes.c_int64 * n)(*[t.numel() for t in tensors])
    kk = (ctypes.c_int64 * n)(*[int(k) for k in ks])
    nz = (ctypes.c_int64 * n)()
    _lib.check(_lib.lib().wt_min_weight_prune_f32(ptrs, ns, kk, n, _stream_ptr(), nz))
    return list(nz)


def scatter_zero_(w: torch.Tensor, idx: torch.Tensor) -> int:
    """w.view(-1)[idx] = 0 in place (w contiguous float32 CUDA, idx int64 CUDA); returns count_nonzero(w)"""
    _check_fp32(w)
    if not w.is_contiguous():
        raise ValueError('scatter_zero_ needs a contiguous tensor')
    idx = idx.to(device=w.device, dtype=torch.int64).contiguous()
    out = torch.empty(1, dtype=torch.int64, device=w.device)
    _lib.check(_lib.lib().wt_scatter_zero_f32(w.data_ptr(), w.numel(), idx.data_ptr(), idx.numel(), out.data_ptr(),
                                              _stream_ptr()))
    return int(out.item())


def scatter_zero_group_(ws: List[torch.Tensor], idxs: List[torch.Tensor]) -> List[int]:
    """scatter_zero_ for a whole model with ONE host synchronisation: every layer's scatter + count is enqueued
    back to back (the counts land in one device tensor), then a single D2H brings all of them back.  The reference
    syncs once per layer (random_pruning.py:57-58)."""
    if not ws:
        return []
    dev = ws[0].device
    counts = torch.empty(len(ws), dtype=torch.int64, device=dev)
    L, s = _lib.lib(), _stream_ptr()
    keep = []
    for i, (w, idx) in enumerate(zip(ws, idxs)):
        _check_fp32(w)
        if not w.is_contiguous():
            raise ValueError('scatter_zero_group_ needs contiguous tensors')
        idx = idx.to(device=dev, dtype=torch.int64).contiguous()
        keep.append(idx)
        _lib.check(L.wt_scatter_zero_f32(w.data_ptr(), w.numel(), idx.data_ptr(), idx.numel(), counts[i:].data_ptr(), s))
    return [int(v) for v in counts.cpu().tolist()]


def scatter_zero_group_async_(ws: List[torch.Tensor], idxs: List[torch.Tensor]) -> torch.Tensor:
    """scatter_zero_group_ without the host synchronisation: returns the int64 CUDA tensor of nonzero counts, valid in
    stream order (the caller reads it when it needs the numbers)"""
    dev = ws[0].device if ws else cuda_device()
    counts = torch.empty(len(ws), dtype=torch.int64, device=dev)
    L, s = _lib.lib(), _stream_ptr()
    for i, (w, idx) in enumerate(zip(ws, idxs)):
        _check_fp32(w)
        if not w.is_contiguous():
            raise ValueError('scatter_zero_group_async_ needs contiguous tensors')
        idx = idx.to(device=dev, dtype=torch.int64).contiguous()
        idx.record_stream(torch.cuda.current_stream())
        _lib.check(L.wt_scatter_zero_f32(w.data_ptr(), w.numel(), idx.data_ptr(), idx.numel(), counts[i:].data_ptr(), s))
    return counts


def count_nonzero_and_below(x: torch.Tensor, eps: float = 1e-6) -> Tuple[int, int]:
    """(torch.count_nonzero(x), (|x| < eps).sum())  -- calculate_sparsity, testing_suite/eval_model.py:7-20"""
    _check_fp32(x)
    x = x.contiguous()
    out = torch.empty(2, dtype=torch.int64, device=x.device)
    _lib.check(_lib.lib().wt_count_f32(x.data_ptr(), x.numel(), float(eps), out.data_ptr(), _stream_ptr()))
    a, b = out.cpu().tolist()
    return int(a), int(b)
