"""cl-waffe2_b200 — B200-native `CUDATensor` device backend for cl-waffe2's data-parallel hot path.

Layout
  csrc/     hand-written sm_100a CUDA + the C ABI (include/waffe2cuda.h) -> lib/libwaffe2cuda.so
  lisp/     the Common Lisp host side a cl-waffe2 maintainer adds as source/backends/cuda/
            (CFFI bindings + define-impl-op registrations); cannot be loaded in this image (no SBCL)
  *.py      the same host logic in Python over ctypes, mirroring the reference's names, so the
            path can be driven, tested and measured here

The directory name is not a Python identifier; import it with
    wf = importlib.import_module("cl-waffe2_b200")
"""
from . import _lib
from ._lib import Wf2CudaError, init, shutdown
from .tensor import (CUDATensor, StaticPool, Storage, T, tensor_at, as_strided, ax_plus_b, calc_strides, from_numpy, make_tensor, permute,
                     range_nth, range_size, reshape_contiguous, transpose, view)
from . import ops

__all__ = ["_lib", "Wf2CudaError", "init", "shutdown", "CUDATensor", "StaticPool", "Storage", "tensor_at", "as_strided", "T", "ax_plus_b", "calc_strides",
           "from_numpy", "make_tensor", "permute", "range_nth", "range_size", "reshape_contiguous", "transpose",
           "view", "ops"]
