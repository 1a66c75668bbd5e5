"""Shim with the reference's flat module name (ResNet/utils.py): put this directory first on sys.path and
ResNet/main_pruning.py:53-55 imports the B200 implementation unchanged."""
from wavelettransforms_b200.utils import *  # noqa: F401,F403
from wavelettransforms_b200 import utils as _impl

globals().update({k: getattr(_impl, k) for k in dir(_impl) if not k.startswith('__')})
