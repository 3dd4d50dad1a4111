"""cufhe_b200 -- B200-native TFHE gate-bootstrapping engine behind cuFHE's API.

csrc/   hand-written sm_100a CUDA kernels + the C ABI (include/b2fhe.h)
api.py  host-side mirror of the reference's `namespace cufhe` interface
"""
from . import _lib  # noqa: F401
from .api import *  # noqa: F401,F403
from .api import OP, OPS, partition  # noqa: F401
