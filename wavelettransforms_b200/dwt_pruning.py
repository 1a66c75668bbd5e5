"""Drop-in for /root/reference/ResNet/dwt_pruning.py (percentile-threshold DWT pruning): same
function names, arguments, return values and aliasing behaviour; the arithmetic runs on the GPU
through libwtprune.  Extra keyword-only arguments never change the positional signatures."""
import os
from typing import List, Optional, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import _lib, config, engine
from .utils import (append_to_experiment_log, check_and_set_pruned_instance_path, log_pruning_details, save_model,
                    setup_csv_writer)


def calculate_max_level(shape, wavelet):
    """dwt_pruning.py:12-13"""
    return engine.calculate_max_level(shape, wavelet)


def analyze_pruning(model: nn.Module):
    """dwt_pruning.py:16-22 -- per-layer sparsity printout (zero counts come from the GPU)"""
    for name, module in model.named_modules():
        if isinstance(module, nn.Conv2d):
            w = module.weight.data
            if w.numel() == 0:
                continue
            x = engine.to_device_f32(w, engine.cuda_device())
            nonzero, _ = engine.count_nonzero_and_below(x)
            print(f"Layer {name}: Sparsity = {(w.numel() - nonzero) / w.numel():.2%}")


def percentile_based_thresholding(coeff_arr, percentile=90, *, numpy_semantics: Optional[str] = None):
    """dwt_pruning.py:25-32: t = np.percentile(|c|, p); where(|c| < t, 0, c).  numpy array in/out."""
    arr = np.asarray(coeff_arr)
    if arr.dtype != np.float32:
        raise NotImplementedError(f'float32 coefficient arrays only (got {arr.dtype})')
    if arr.size == 0:
        raise IndexError('index -1 is out of bounds for axis 0 with size 0')  # what np.percentile raises
    x = torch.from_numpy(np.ascontiguousarray(arr)).to(engine.cuda_device())
    res = engine.prune_device_tensors([x.view(-1)], [0], 'haar', 'periodization', _lib.THRESHOLD_PERCENTILE,
                                      float(percentile), numpy_semantics)
    if config.verbose:
        print(f"Percentile: {percentile}, Threshold: {res.thresholds[0]}")
    return res.outputs[0].view(arr.shape).cpu().numpy()


def _effective_levels(shapes, wavelet, level, clamp_level=True):
    """dwt_pruning.py:64-65 incl. the sticky rebinding of `level` across the list"""
    out = []
    for shp in shapes:
        if len(shp) < 2:
            out.append(0)
            continue
        if clamp_level:
            level = min(level, calculate_max_level(shp, wavelet))
        out.append(level)
    return out


def multi_resolution_analysis(weights: List[torch.Tensor], wavelet: str, level: int, percentile: float,
                              mode: str = 'periodization', *, numpy_semantics: Optional[str] = None,
                              clamp_level: bool = True, arithmetic: Optional[str] = None) -> Tuple[List[torch.Tensor], int]:
    """dwt_pruning.py:35-95.  Returns new tensors (dtype/device/shape of the inputs, inputs untouched)
    and the total number of zeros in them (torch.sum(pruned == 0), :88)."""
    _lib.wavelet_lookup(wavelet)  # ValueError for unknown names, as pywt.Wavelet does
    _lib.mode_index(mode)
    if level < 0:
        raise ValueError(f"Level value of {level} is too low . Minimum level is 0.")
    weights = list(weights)
    if not weights:
        return [], 0
    dev = engine.cuda_device()
    levels = _effective_levels([tuple(w.shape) for w in weights], wavelet, level, clamp_level)
    xs = [engine.to_device_f32(w, dev) for w in weights]
    live = [i for i, x in enumerate(xs) if x.numel() > 0]
    if len(live) != len(xs):
        raise IndexError('index -1 is out of bounds for axis 0 with size 0')  # np.percentile on an empty array
    res = engine.prune_device_tensors(xs, levels, wavelet, mode, _lib.THRESHOLD_PERCENTILE, float(percentile),
                                      numpy_semantics, arithmetic=arithmetic)
    pruned = [o.view(w.shape).to(w.device) for o, w in zip(res.outputs, weights)]
    return pruned, int(res.counts[:, 2].sum())


def prune_layer_weights(layer: nn.Module, wavelet: str, level: int, percentile: float) -> Tuple[int, int, int]:
    """dwt_pruning.py:98-127: (numel, nonzeros after, zeros after); rebinds layer.weight.data"""
    with torch.no_grad():
        weights = [layer.weight]
        original_param_count = sum(p.numel() for p in weights)
        pruned_weights, total_pruned_count = multi_resolution_analysis(weights, wavelet, level, percentile)
        non_zero_params = original_param_count - total_pruned_count
        if config.verbose:
            print(f"Original Param Count: {original_param_count}, Non-zero Params: {non_zero_params}, "
                  f"Total Pruned Count: {total_pruned_count}")
        layer.weight.data = pruned_weights[0]
        return original_param_count, non_zero_params, total_pruned_count


def prune_model_layers(model: nn.Module, wavelet: str, level: int, percentile: float, *, mode: str = 'periodization',
                       numpy_semantics: Optional[str] = None, clamp_level: bool = True):
    """All nn.Conv2d weights of `model` in ONE grouped plan (the per-layer loop of dwt_pruning.py:155-164
    without its per-layer host round trips).  Returns [(name, numel, nonzero, zeros)] in named_modules()
    order; weights are rebound like prune_layer_weights does."""
    layers = [(n, m) for n, m in model.named_modules() if isinstance(m, nn.Conv2d)]
    if not layers:
        return []
    _lib.wavelet_lookup(wavelet)
    if level < 0:
        raise ValueError(f"Level value of {level} is too low . Minimum level is 0.")
    dev = engine.cuda_device()
    with torch.no_grad():
        ws = [m.weight for _, m in layers]
        # each layer is its own multi_resolution_analysis call in the reference: no sticky clamp across layers
        levels = [_effective_levels([tuple(w.shape)], wavelet, level, clamp_level)[0] for w in ws]
        xs = [engine.to_device_f32(w, dev) for w in ws]
        res = engine.prune_device_tensors(xs, levels, wavelet, mode, _lib.THRESHOLD_PERCENTILE, float(percentile),
                                          numpy_semantics)
        rows = []
        for (name, m), o, c in zip(layers, res.outputs, res.counts):
            m.weight.data = o.view(m.weight.shape).to(m.weight.device)
            rows.append((name, int(c[0]), int(c[1]), int(c[2])))
    return rows


def wavelet_pruning(model, wavelet: str, level: int, percentile: float, csv_path: str, guid: str) -> str:
    """dwt_pruning.py:130-174: prune every Conv2d, write the per-layer log, save the model, append the
    experiment row; returns the per-layer log path."""
    threshold_value = percentile / 100
    selective_pruned_dir = check_and_set_pruned_instance_path(
        f"{wavelet}_threshold-{threshold_value}_level-{level}_guid-{guid[:4]}/selective_pruned")
    selective_log_path = os.path.join(selective_pruned_dir, 'log.csv')
    writer, log_file = setup_csv_writer(os.path.normpath(selective_log_path), mode='w')
    total_pruned_count = 0
    total_non_zero_params = 0
    try:
        for name, numel, nonzero, zeros in prune_model_layers(model, wavelet, level, percentile):
            total_pruned_count += zeros
            total_non_zero_params += nonzero
            log_pruning_details(writer, guid, wavelet, level, threshold_value, 'selective', numel, nonzero, zeros, name)
    finally:
        log_file.close()
    save_model(model, selective_pruned_dir)
    append_to_experiment_log(os.path.normpath(csv_path), guid, wavelet, level, threshold_value, 'selective',
                             total_pruned_count, total_non_zero_params, selective_pruned_dir)
    print(f"Selectively pruned model saved at {selective_pruned_dir}")
    return selective_log_path
