"""fempar_b200: B200-native (sm_100a CUDA + NCCL) implementation of FEMPAR's assembly + PCG hot path.

Only the hot path lives here: csrc/ (CUDA kernels + C-ABI, built into libfempar_b200.so) and the
host-side mirror of the reference interface (fempar.py).  No CPU fallback exists.
"""
from . import _lib
from ._lib import Fb2Error, init, load
from .fempar import *  # noqa: F401,F403
