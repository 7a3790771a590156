"""photon_ml_b200 -- B200-native GLM objective evaluation behind Photon ML's
ObjectiveFunction / DiffFunction / TwiceDiffFunction interface.

The product is the CUDA library photon_ml_b200/lib/libphoton_ml_b200.so (C ABI in
include/photon_ml_b200.h).  This package is the thin host side used by tests and bench.py: a
ctypes binding (abi) and a mirror of the reference's objective-function classes (function).
"""
from . import abi  # noqa: F401
from .abi import (Batch, Context, Dataset, IllegalArgumentError, Peer, PhotonError, LOGISTIC, POISSON, SQUARED,  # noqa: F401
                  SMOOTHED_HINGE, STORE_AUTO, STORE_F32, STORE_F64, STORE_ONES)
