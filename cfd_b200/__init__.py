"""cfd_b200 -- B200-native (sm_100a) backend for the shaia/CFD projection-method time step.

Host-side mirror of the reference's solver interface over the C-ABI of libcfd_b200.so
(include/cfd_b200.h).  There is no CPU fallback: anything that computes raises if the CUDA
library is missing.
"""
from . import abi  # noqa: F401

__version__ = "0.1.0"
