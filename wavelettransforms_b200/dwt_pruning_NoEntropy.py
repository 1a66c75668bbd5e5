"""Drop-in for /root/reference/ResNet/dwt_pruning_NoEntropy.py (absolute-threshold DWT pruning, no
level clamp).  Same names, arguments, return values and aliasing; arithmetic on the GPU."""
import os
from typing import List, Tuple

import torch
import torch.nn as nn

from . import _lib, engine
from .utils import (append_to_experiment_log, check_and_set_pruned_instance_path, log_pruning_details, save_model,
                    setup_csv_writer)


def multi_resolution_analysis(weights: List[torch.Tensor], wavelet: str, level: int, threshold: float,
                              mode: str = 'periodization', *, arithmetic=None) -> Tuple[List[torch.Tensor], int]:
    """dwt_pruning_NoEntropy.py:12-62: wavedec2(level) with no clamp, coeff_arr[|c| < threshold] = 0,
    waverec2, crop.  Returns (new tensors, sum over tensors of nonzeros(before) - nonzeros(after))."""
    _lib.wavelet_lookup(wavelet)
    _lib.mode_index(mode)
    if level < 0:
        raise ValueError(f"Level value of {level} is too low . Minimum level is 0.")
    weights = list(weights)
    if not weights:
        return [], 0
    dev = engine.cuda_device()
    xs = [engine.to_device_f32(w, dev) for w in weights]
    levels = [level if w.ndim >= 2 else 0 for w in weights]
    res = engine.prune_device_tensors(xs, levels, wavelet, mode, _lib.THRESHOLD_ABSOLUTE, float(threshold), arithmetic=arithmetic)
    pruned = [o.view(w.shape).to(w.device) for o, w in zip(res.outputs, weights)]
    total = int((res.nonzero_before - res.counts[:, 1]).sum())
    return pruned, total


def prune_layer_weights(layer: nn.Module, wavelet: str, level: int, threshold: float) -> Tuple[int, int, int]:
    """dwt_pruning_NoEntropy.py:65-91: all parameters of the layer; returns (numel, nonzeros BEFORE
    pruning, pruned count); writes back with param.copy_ when requires_grad"""
    weights = [p for p in layer.parameters()]
    original_param_count = sum(w.numel() for w in weights)
    if not weights:
        return 0, 0, 0
    dev = engine.cuda_device()
    xs = [engine.to_device_f32(w, dev) for w in weights]
    levels = [level if w.ndim >= 2 else 0 for w in weights]
    _lib.wavelet_lookup(wavelet)
    if level < 0:
        raise ValueError(f"Level value of {level} is too low . Minimum level is 0.")
    res = engine.prune_device_tensors(xs, levels, wavelet, 'periodization', _lib.THRESHOLD_ABSOLUTE, float(threshold))
    non_zero_params = int(res.nonzero_before.sum())
    layer_pruned_count = int((res.nonzero_before - res.counts[:, 1]).sum())
    with torch.no_grad():
        for param, o in zip(layer.parameters(), res.outputs):
            if param.requires_grad:
                param.copy_(o.view(param.shape).to(param.device))
    return original_param_count, non_zero_params, layer_pruned_count


def wavelet_pruning(model, wavelet: str, level: int, threshold: float, csv_path: str, guid: str) -> str:
    """dwt_pruning_NoEntropy.py:94-138"""
    selective_pruned_dir = check_and_set_pruned_instance_path(
        f"{wavelet}_threshold-{threshold}_level-{level}_guid-{guid[:4]}/selective_pruned")
    selective_log_path = os.path.join(selective_pruned_dir, 'log.csv')
    writer, log_file = setup_csv_writer(os.path.normpath(selective_log_path), mode='w')
    total_pruned_count = 0
    total_non_zero_params = 0
    try:
        # ONE plan for the whole model (the layers are independent: dwt_pruning_NoEntropy.py:79-89 handles them one by
        # one); per-layer bookkeeping afterwards from the counts the kernels left
        _lib.wavelet_lookup(wavelet)
        if level < 0:
            raise ValueError(f"Level value of {level} is too low . Minimum level is 0.")
        layers = [(name, module) for name, module in model.named_modules() if isinstance(module, nn.Conv2d)]
        params = [[p for p in module.parameters()] for _, module in layers]
        flat_params = [p for ps in params for p in ps]
        if flat_params:
            dev = engine.cuda_device()
            xs = [engine.to_device_f32(p, dev) for p in flat_params]
            levels = [level if p.ndim >= 2 else 0 for p in flat_params]
            res = engine.prune_device_tensors(xs, levels, wavelet, 'periodization', _lib.THRESHOLD_ABSOLUTE, float(threshold))
        at = 0
        for (name, module), ps in zip(layers, params):
            numel = sum(p.numel() for p in ps)
            sl = slice(at, at + len(ps))
            non_zero = int(res.nonzero_before[sl].sum()) if ps else 0
            pruned = int((res.nonzero_before[sl] - res.counts[sl, 1]).sum()) if ps else 0
            with torch.no_grad():
                for param, o in zip(ps, res.outputs[sl] if ps else []):
                    if param.requires_grad:
                        param.copy_(o.view(param.shape).to(param.device))
            at += len(ps)
            total_pruned_count += pruned
            total_non_zero_params += non_zero
            log_pruning_details(writer, guid, wavelet, level, threshold, 'selective', numel, non_zero, pruned, name)
    finally:
        log_file.close()
    save_model(model, selective_pruned_dir)
    append_to_experiment_log(os.path.normpath(csv_path), guid, wavelet, level, threshold, 'selective',
                             total_pruned_count, total_non_zero_params, selective_pruned_dir)
    print(f"Selectively pruned model saved at {selective_pruned_dir}")
    return selective_log_path
