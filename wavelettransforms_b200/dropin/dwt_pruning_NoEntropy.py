"""Shim with the reference's flat module name (ResNet/dwt_pruning_NoEntropy.py): put this directory first on sys.path and
ResNet/main_pruning.py:53-55 imports the B200 implementation unchanged."""
from wavelettransforms_b200.dwt_pruning_NoEntropy import *  # noqa: F401,F403
from wavelettransforms_b200 import dwt_pruning_NoEntropy as _impl

globals().update({k: getattr(_impl, k) for k in dir(_impl) if not k.startswith('__')})
