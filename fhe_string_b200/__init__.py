"""fhe_string_b200 -- B200-native batched KS-PBS engine behind the fhe_string API.

The product is the C-ABI shared library `libfhs_cuda.so` (include/fhs_cuda.h, sources under
fhe_string_b200/csrc).  This package is the thin Python host binding used by tests and bench.py.
There is no CPU fallback: importing is cheap, but creating an Engine without the built library or
without a CUDA device raises.
"""
from .engine import Engine, FhsError, Params, lib_path, load_library  # noqa: F401

__all__ = ["Engine", "FhsError", "Params", "lib_path", "load_library"]
