"""Drop-in for /root/reference/ResNet/random_pruning.py: per layer, zero as many randomly chosen
weights as the DWT run pruned there.  torch.randperm stays the RNG (same generator state => same
permutation => bit-identical mask); the scatter and the nonzero count run on the GPU."""
import csv
import os
from queue import Queue
from typing import Optional

import torch
import torch.nn as nn

from . import engine
from .utils import (append_to_experiment_log, check_and_set_pruned_instance_path, get_layer, log_pruning_details,
                    save_model, setup_csv_writer)


def random_prune_(weights: torch.Tensor, prune_count: int) -> int:
    """random_pruning.py:51-58 for one tensor, in place; returns count_nonzero afterwards.
    The permutation is drawn exactly as the reference draws it: torch.randperm(n, device=weights.device)."""
    flat = weights.view(-1)
    idx = torch.randperm(flat.numel(), device=weights.device)[:prune_count]
    if weights.is_cuda:
        return engine.scatter_zero_(flat, idx)
    dev = engine.cuda_device()
    work = engine.to_device_f32(flat, dev)
    nz = engine.scatter_zero_(work, idx.to(dev))
    flat.copy_(work)
    return nz


def random_pruning(model, selective_log_path: str, guid: str, wavelet: str, level: int, threshold: float,
                   csv_path: str, log_queue: Optional[Queue] = None) -> None:
    """random_pruning.py:11-82"""
    total_pruned_count = 0
    total_non_zero_params = 0
    random_pruned_dir = check_and_set_pruned_instance_path(
        f"{wavelet}_threshold-{threshold}_level-{level}_guid-{guid[:4]}/random_pruned")
    random_log_path = os.path.join(random_pruned_dir, 'log.csv')
    writer, log_file = setup_csv_writer(os.path.normpath(random_log_path), mode='w')
    # One pass over the selective log collects the layers; the permutations are drawn in log order exactly as the
    # reference draws them (one torch.randperm(n, device=weights.device) per layer, random_pruning.py:53-54), then every
    # layer's scatter + count is enqueued back to back and ONE device-to-host copy brings the counts (the reference
    # synchronises per layer, :57-58).
    entries = []
    with open(selective_log_path, 'r') as f:
        for row in csv.DictReader(f):
            layer_name = row['Layer Name']
            layer = get_layer(model, layer_name)
            if layer is not None and isinstance(layer, nn.Conv2d):
                entries.append((layer_name, layer, int(row['Original Parameter Count']), int(row['Total Pruned Count'])))
            else:
                print(f"Layer not found or not a Conv2D layer: {layer_name}")
    with torch.no_grad():
        dev = engine.cuda_device()
        work, idxs, homes = [], [], []
        for _, layer, _, prune_count in entries:
            weights = layer.weight.data
            flat = weights.view(-1)
            idx = torch.randperm(flat.numel(), device=weights.device)[:prune_count]
            if weights.is_cuda:
                work.append(flat)
                homes.append(None)
            else:
                work.append(engine.to_device_f32(flat, dev))
                homes.append(flat)
            idxs.append(idx)
        nonzeros = engine.scatter_zero_group_(work, idxs)
        for (_, layer, _, _), w, home in zip(entries, work, homes):
            if home is not None:
                home.copy_(w)
            layer.weight.data = layer.weight.data.view(-1).view_as(layer.weight.data)
    for (layer_name, _, original_param_count, _), nonzero in zip(entries, nonzeros):
        actual_pruned = original_param_count - nonzero
        log_pruning_details(writer, guid, wavelet, level, threshold, 'random', original_param_count, nonzero,
                            actual_pruned, layer_name)
        total_pruned_count += actual_pruned
        total_non_zero_params += nonzero
    save_model(model, random_pruned_dir)
    if log_queue is not None:
        log_queue.put((guid, wavelet, level, threshold, 'random', total_pruned_count, total_non_zero_params,
                       random_pruned_dir))
    else:
        append_to_experiment_log(csv_path, guid, wavelet, level, threshold, 'random', total_pruned_count,
                                 total_non_zero_params, random_pruned_dir)
    log_file.close()
    print("Random pruning completed.")
