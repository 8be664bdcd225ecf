"""paffy_b200 - B200-native implementation of paffy's add_mismatches hot path.

The product is the CUDA library libpaffy_b200.so (C ABI in include/paffy_b200.h); this package is
the thin Python plumbing over it (ctypes), mirroring the reference's interface for the path.
"""
from .lib import Batch, Context, PafGpuError, Result, load, ABI_SYMBOLS  # noqa: F401

__all__ = ["Context", "PafGpuError", "Batch", "Result", "load", "ABI_SYMBOLS"]
