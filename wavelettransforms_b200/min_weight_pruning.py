"""Drop-in for /root/reference/ResNet/min_weight_pruning.py: smallest-magnitude masks at the overall
sparsity the DWT run reached.  Same names / arguments / in-place behaviour; the exact-k select, the
zeroing and the nonzero count run on the GPU (libwtprune wt_min_weight_prune_f32)."""
import csv
import os
from queue import Queue
from typing import Dict, Optional

import torch

from . import engine
from .utils import (append_to_experiment_log, check_and_set_pruned_instance_path, log_pruning_details, save_model,
                    setup_csv_writer)


def read_selective_pruning_log(selective_log_path: str) -> Dict[str, int]:
    """min_weight_pruning.py:42-51"""
    pruned_layers = {}
    with open(selective_log_path, 'r') as f:
        for row in csv.DictReader(f):
            pruned_layers[row['Layer Name']] = int(row['Original Parameter Count'])
    return pruned_layers


def calculate_dwt_pruning_percentage(selective_log_path: str) -> float:
    """min_weight_pruning.py:54-63"""
    total_params = 0
    total_pruned = 0
    with open(selective_log_path, 'r') as f:
        for row in csv.DictReader(f):
            total_params += int(row['Original Parameter Count'])
            total_pruned += int(row['Total Pruned Count'])
    return total_pruned / total_params if total_params > 0 else 0.0


def _prune_inplace(tensors, prune_percentage):
    """shared by percentage_min_pruning and the walker: mutates each tensor's storage, returns nonzero counts"""
    dev = engine.cuda_device()
    ks = [int(t.numel() * prune_percentage) for t in tensors]   # :70, float64 product then truncation
    for t, k in zip(tensors, ks):
        if k > t.numel() or k < 0:
            raise RuntimeError('selected index k out of range')  # torch.topk's error
    work = []
    for t in tensors:
        if not t.is_contiguous():
            raise RuntimeError('view size is not compatible with input tensor\'s size and stride')  # .view(-1), :68
        work.append(t if t.is_cuda else engine.to_device_f32(t, dev))
    nz = engine.min_weight_prune_(work, ks)
    for t, wk in zip(tensors, work):
        if wk is not t:
            t.copy_(wk)            # same storage as the caller's tensor, like flatten_weights[idx] = 0
    return nz


def percentage_min_pruning(weights: torch.Tensor, prune_percentage: float) -> torch.Tensor:
    """min_weight_pruning.py:66-74: zero the int(n*pct) smallest |w| IN PLACE; returns a view of `weights`.
    Ties at the k-th magnitude: lowest flat index first (torch.topk leaves this unspecified)."""
    with torch.no_grad():
        _prune_inplace([weights], prune_percentage)
    return weights.view(-1).view_as(weights)


def min_weight_pruning(model, selective_log_path: str, guid: str, wavelet: str, level: int, threshold: float,
                       csv_path: str, log_queue: Optional[Queue] = None) -> None:
    """min_weight_pruning.py:77-139"""
    print(f"Starting min_weight_pruning with GUID: {guid}")
    min_pruned_dir = check_and_set_pruned_instance_path(
        f"{wavelet}_threshold-{threshold}_level-{level}_guid-{guid[:4]}/min_pruned")
    min_log_path = os.path.join(min_pruned_dir, 'log.csv')
    writer, log_file = setup_csv_writer(min_log_path, mode='w')
    try:
        pct = calculate_dwt_pruning_percentage(selective_log_path)
        print(f"Overall pruning percentage from DWT: {pct:.2%}")
        pruned_layers = read_selective_pruning_log(selective_log_path)
        targets = []
        for name, module in model.named_modules():
            if name in pruned_layers and hasattr(module, 'weight') and isinstance(module.weight, torch.Tensor):
                targets.append((name, module))
        total_params = 0
        total_pruned = 0
        with torch.no_grad():
            nz = _prune_inplace([m.weight.data for _, m in targets], pct) if targets else []
        for (name, module), nonzero in zip(targets, nz):
            original = pruned_layers[name]
            actual_pruned = original - nonzero
            log_pruning_details(writer, guid, wavelet, level, threshold, 'min', original, nonzero, actual_pruned, name)
            total_params += original
            total_pruned += actual_pruned
        save_model(model, min_pruned_dir)
        if log_queue is not None:
            log_queue.put((guid, wavelet, level, threshold, 'min', total_pruned, total_params - total_pruned,
                           min_pruned_dir))
        else:
            append_to_experiment_log(csv_path, guid, wavelet, level, threshold, 'min', total_pruned,
                                     total_params - total_pruned, min_pruned_dir)
        log_file.close()
        print(f"Minimum weight pruning completed. Pruned {total_pruned} out of {total_params} parameters.")
        if total_params:
            print(f"Actual pruning percentage: {total_pruned/total_params:.2%}")
    except Exception as e:
        print(f"Error in min_weight_pruning: {str(e)}")
        raise
