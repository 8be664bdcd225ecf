"""paffy_b200 - B200-native implementation of paffy's add_mismatches hot path.

The product is the CUDA library libpaffy_b200.so (C ABI in include/paffy_b200.h); this package is
the thin Python plumbing over it (ctypes), mirroring the reference's interface for the path.
"""
from .lib import (ABI_SYMBOLS, REFERENCE_SYMBOLS, Batch, Context, FileOpts, FileReport, PafGpuError, Result, load, partition,  # noqa: F401
                  remove_mismatches_file)

__all__ = ["Context", "PafGpuError", "Batch", "Result", "FileOpts", "FileReport", "load", "partition", "remove_mismatches_file", "ABI_SYMBOLS",
           "REFERENCE_SYMBOLS"]
