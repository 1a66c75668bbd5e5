"""The three pruning methods of one experiment on three CUDA streams.

The reference's driver (ResNet/main_pruning.py:175-209) deep-copies the model three times, runs the DWT pruner on the main
thread, and then starts two Python threads for the comparison maskers, which prune as many weights as the DWT did: the
smallest-magnitude masker takes the DWT's overall pruned fraction (min_weight_pruning.py:25-33, :66-74), the random masker
the DWT's per-layer pruned count (random_pruning.py:40-58).  Here the weights are uploaded once, the DWT prune is one plan on
the caller's stream, and the two maskers are enqueued on two more streams that fork from it -- the random masker first,
because it never needs the host, then the smallest-magnitude masker, whose grouped entry point returns its counts.  The
arithmetic is that of the three drop-in walkers (same kernels, same torch.randperm draws in the same order)."""
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib, engine
from .dwt_pruning import _effective_levels


def prune_three_ways(weights: Sequence[torch.Tensor], wavelet: str, level: int, percentile: float, *,
                     mode: str = 'periodization', clamp_level: bool = True, numpy_semantics: Optional[str] = None,
                     arithmetic: Optional[str] = None) -> Dict[str, Tuple[List[torch.Tensor], np.ndarray]]:
    """weights: float32 tensors (host or device).  Returns {'selective' | 'min' | 'random': (pruned CUDA tensors,
    int64 [n, 3] of (numel, nonzero, zeros) per tensor)}; the input tensors are not modified."""
    dev = engine.cuda_device()
    main = torch.cuda.current_stream()
    xs = [engine.to_device_f32(w, dev) for w in weights]
    if not xs:
        empty = np.zeros((0, 3), dtype=np.int64)
        return {'selective': ([], empty), 'min': ([], empty), 'random': ([], empty)}
    levels = _effective_levels([tuple(x.shape) for x in xs], wavelet, level, clamp_level)
    res = engine.prune_device_tensors(xs, levels, wavelet, mode, _lib.THRESHOLD_PERCENTILE, float(percentile), numpy_semantics,
                                      arithmetic=arithmetic)
    numel = res.counts[:, 0]
    pruned = numel - res.counts[:, 1]                         # 'Total Pruned Count' of the selective log, per layer
    total = int(numel.sum())
    pct = float(pruned.sum()) / total if total else 0.0       # calculate_dwt_pruning_percentage
    s_min, s_rnd = torch.cuda.Stream(), torch.cuda.Stream()
    s_min.wait_stream(main)
    s_rnd.wait_stream(main)
    with torch.no_grad():
        with torch.cuda.stream(s_rnd):
            for x in xs:
                x.record_stream(s_rnd)
            rnd = [x.clone() for x in xs]
            idxs = [torch.randperm(r.numel(), device=dev)[:int(k)] for r, k in zip(rnd, pruned)]   # random_pruning.py:53-54
            rnd_counts = engine.scatter_zero_group_async_([r.view(-1) for r in rnd], idxs)
        with torch.cuda.stream(s_min):
            for x in xs:
                x.record_stream(s_min)
            mn = [x.clone() for x in xs]
            ks = [int(m.numel() * pct) for m in mn]                                               # min_weight_pruning.py:70
            mn_nz = engine.min_weight_prune_(mn, ks)          # returns when stream s_min is done; s_rnd keeps running
    main.wait_stream(s_rnd)                                   # the caller's later work is ordered behind both maskers
    main.wait_stream(s_min)
    s_rnd.synchronize()
    rnd_nz = rnd_counts.cpu().numpy()
    for t in rnd + mn:
        t.record_stream(main)

    def table(nz):
        nz = np.asarray(nz, dtype=np.int64)
        return np.stack([numel, nz, numel - nz], axis=1)
    return {'selective': (res.outputs, res.counts), 'min': (mn, table(mn_nz)), 'random': (rnd, table(rnd_nz))}
