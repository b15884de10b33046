"""mundy_b200: B200-native (sm_100a) collision hot path of MuNDy behind a C-ABI.

The product is mundy_b200/libmundy_b200.so (include/mundy_b200.h). This package only binds it:
  _capi   ctypes declarations of every exported symbol (import fails loudly if the library is not built)
  ops     torch-tensor convenience wrappers (device memory and streams come from torch)
  workloads  synthetic body generators for the BASELINE.json configurations
"""
from . import _capi  # noqa: F401  (raises ImportError when the CUDA library is missing: no CPU fallback)

__all__ = ["_capi"]
