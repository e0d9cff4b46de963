"""trajopt-1_b200: B200-native collision-term evaluation for trajopt (one hot path).

The product is the C-ABI library csrc/libtrajopt_cuda.so (include/trajopt_cuda.h) plus the C++ host
adaptors in host/ that mirror the reference's evaluator classes.  This Python package only contains
the ctypes driver used by tests and benchmarks, the host-side multi-start logic (multistart.py) and the batched
mirrors of the callers next to the evaluators (constraints.py: numerical-gradient constraints, trajectory check)."""
from . import capi, constraints, multistart  # noqa: F401
from .capi import CudaLib, LIB_PATH  # noqa: F401
