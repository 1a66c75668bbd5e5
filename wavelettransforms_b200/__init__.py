"""wavelettransforms_b200 -- B200 (sm_100a) implementation of the DWT-prune hot path of
iAmGiG/WaveletTransforms and its two sibling maskers, behind the reference's own module and
function names:

    wavelettransforms_b200.dwt_pruning            <- ResNet/dwt_pruning.py
    wavelettransforms_b200.dwt_pruning_NoEntropy  <- ResNet/dwt_pruning_NoEntropy.py
    wavelettransforms_b200.min_weight_pruning     <- ResNet/min_weight_pruning.py
    wavelettransforms_b200.random_pruning         <- ResNet/random_pruning.py

The arithmetic runs in hand-written CUDA kernels reached through the C ABI of
lib/libwtprune.so (include/wtprune.h).  There is no CPU fallback.
"""
from . import config  # noqa: F401

__all__ = ['config']
__version__ = '0.1.0'
