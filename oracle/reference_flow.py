"""oracle/reference_flow.py -- TEST INFRASTRUCTURE ONLY.

Line-for-line CPU restatement of the reference's per-tensor flows, with the oracle's C functions
standing in for the four pywt calls and NumPy / PyTorch used exactly where the reference uses
them.  This is also the timed CPU baseline (bench.py cpu_baseline / --impl reference).

  percentile flow : /root/reference/ResNet/dwt_pruning.py:25-32, 35-95
  absolute flow   : /root/reference/ResNet/dwt_pruning_NoEntropy.py:12-62
  min-weight      : /root/reference/ResNet/min_weight_pruning.py:66-74
  random          : /root/reference/ResNet/random_pruning.py:51-58

numpy_semantics:
  'live'        -- call np.percentile of the installed NumPy (2.x here => NEP 50 float32 lerp)
  'nep50_2x'    -- restated NumPy 2.x arithmetic (must equal 'live' on this image)
  'legacy_1_22' -- restated arithmetic of the reference's pinned NumPy 1.22.4
"""
import numpy as np
import torch

import oracle as _o

SEM = {'legacy_1_22': 0, 'nep50_2x': 1}


def calculate_max_level(shape, wavelet):
    """dwt_pruning.py:12-13"""
    return _o.dwt_max_level(min(shape[-2:]), _o.dec_len(wavelet))


def percentile_threshold_value(abs_arr, percentile, numpy_semantics='live'):
    """np.percentile(np.abs(arr), percentile) as the float32 the `<` compare sees."""
    if numpy_semantics == 'live':
        return np.float32(np.percentile(abs_arr, percentile))
    sem = SEM[numpy_semantics]
    flat = abs_arr.ravel()
    n = flat.size
    k0, k1, gamma = _o.percentile_plan(n, percentile, sem)
    part = np.partition(flat, sorted({k0, k1}))
    return _o.percentile_lerp(part[k0], part[k1], gamma, sem)


def percentile_based_thresholding(coeff_arr, percentile=90, numpy_semantics='live'):
    """dwt_pruning.py:25-32 (without the debug print)"""
    a = np.abs(coeff_arr)
    threshold = percentile_threshold_value(a, percentile, numpy_semantics)
    return np.where(a < threshold, 0, coeff_arr).astype(coeff_arr.dtype, copy=False)


def multi_resolution_analysis(weights, wavelet, level, percentile, mode='periodization',
                              numpy_semantics='live', clamp_level=True):
    """dwt_pruning.py:35-95.  clamp_level=False skips :64-65 (the NoEntropy decomposition with a
    percentile threshold; not a reference code path, used for the unclamped sweep)."""
    pruned_weights, total = [], 0
    for weight in weights:
        original_shape = weight.shape
        w = weight.detach().cpu().numpy()
        if w.ndim < 2:
            out = percentile_based_thresholding(w, percentile, numpy_semantics)
        else:
            if clamp_level:
                level = min(level, calculate_max_level(w.shape, wavelet))   # sticky, as in :65
            arr = _o.wavedec2_packed(w, wavelet, level, mode)
            arr = percentile_based_thresholding(arr, percentile, numpy_semantics)
            out = _o.waverec2_packed(arr, w.shape, wavelet, level, mode)
            if out.shape != tuple(original_shape):
                out = out[tuple(slice(0, d) for d in original_shape)]
        t = torch.tensor(np.ascontiguousarray(out), dtype=weight.dtype, device=weight.device).view(original_shape)
        pruned_weights.append(t)
        total += int(torch.sum(t == 0).item())
    return pruned_weights, total


def multi_resolution_analysis_absolute(weights, wavelet, level, threshold, mode='periodization'):
    """dwt_pruning_NoEntropy.py:12-62"""
    pruned_weights, total = [], 0
    for weight in weights:
        original_shape = weight.shape
        w = weight.detach().cpu().numpy()
        if w.ndim < 2:
            out = np.where(np.abs(w) < threshold, 0, w).astype(w.dtype)
        else:
            arr = _o.wavedec2_packed(w, wavelet, level, mode)
            arr[np.abs(arr) < threshold] = 0
            out = _o.waverec2_packed(arr, w.shape, wavelet, level, mode)
            out = out[tuple(slice(0, d) for d in original_shape)]
        t = torch.tensor(np.ascontiguousarray(out), dtype=weight.dtype, device=weight.device)
        pruned_weights.append(t)
        total += int(torch.count_nonzero(weight).item()) - int(torch.count_nonzero(t).item())
    return pruned_weights, total


def percentage_min_pruning(weights, prune_percentage):
    """min_weight_pruning.py:66-74 (verbatim semantics, live torch.topk; mutates in place)"""
    flat = weights.view(-1)
    k = int(flat.numel() * prune_percentage)
    _, idx = torch.topk(torch.abs(flat), k, largest=False)
    flat[idx] = 0
    return flat.view_as(weights)


def random_prune_(weights, prune_count):
    """random_pruning.py:51-58 (live torch.randperm on the global generator; in place)"""
    flat = weights.view(-1)
    idx = torch.randperm(flat.numel(), device=weights.device)[:prune_count]
    flat[idx] = 0
    return flat.view_as(weights), int(torch.count_nonzero(flat).item())
