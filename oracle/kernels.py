"""oracle/kernels.py — TEST INFRASTRUCTURE, NOT PRODUCT.

ctypes access to the CPU leaf kernels the oracle drives:

* ``kind == "reference"``: ``oracle/_ref/libwaffe2simd_ref.so`` — the reference's own
  ``cl-waffe2-simd.c`` + vendored SLEEF, compiled by ``oracle/Makefile`` (see its header).
* ``kind == "port"``: ``oracle/_build/libwaffe2_port.so`` — ``oracle/port.c``, our scalar
  restatement with the same ABI.

Calling convention follows the reference's CFFI bindings
(``/root/reference/cl-waffe2-simd/api.lisp:63-274``): ``(n, x, incx, y, incy)`` with element
pointers already displaced by the view offset, as ``incf-tensor-ptr`` does
(``source/backends/cpu/tensor.lisp:31-40``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / reference legs may
import this module.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libwaffe2simd_ref.so")
PORT_SO = os.path.join(HERE, "_build", "libwaffe2_port.so")

_c_long = ctypes.c_long
_c_float = ctypes.c_float
_vp = ctypes.c_void_p

UNARY = ["sin", "cos", "tan", "asin", "acos", "atan", "sinh", "cosh", "tanh",
         "log", "log1p", "log2", "log10", "exp", "abs", "sqrt", "cbrt"]
ARITH = ["add", "sub", "mul", "div"]
CMP = ["eq", "lt", "le", "gt", "ge"]


def build(verbose: bool = False) -> None:
    """Compile the port, and the reference build when /root/reference is present."""
    r = subprocess.run(["make", "-s", "-C", HERE, "all"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout, r.stderr)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)


class Kernels:
    """One loaded leaf-kernel library."""

    def __init__(self, path: str, kind: str):
        self.path = path
        self.kind = kind
        self.lib = ctypes.CDLL(path)
        L = self.lib
        for name in ARITH + UNARY:
            fn = getattr(L, "waffe2_s" + name)
            fn.restype = None
            fn.argtypes = [_c_long, _vp, _c_long, _vp, _c_long]
        L.waffe2_spow.restype = None
        L.waffe2_spow.argtypes = [_c_long, _vp, _c_long, _c_float, _vp, _c_long]
        for name in ("max", "min"):
            fn = getattr(L, "waffe2_s" + name)
            fn.restype = None
            fn.argtypes = [_c_long, _vp, _c_long, _vp]
        for name in CMP:
            fn = getattr(L, "waffe2_s" + name)
            fn.restype = None
            fn.argtypes = [_c_long, _vp, _c_long, _vp, _c_long, _vp, _c_long, _c_float, _c_float]
            fn = getattr(L, "waffe2_s" + name + "_scal")
            fn.restype = None
            fn.argtypes = [_c_long, _vp, _c_long, _vp, _c_long, _c_float, _c_float, _c_float]

        u64 = ctypes.c_uint64
        for name in ("im2col", "col2im"):
            fn = getattr(L, f"waffe2_{name}_s_column")      # the _row variants write out of bounds (port.c note)
            fn.restype = None
            fn.argtypes = [_vp] + [u64] * 14 + [_vp]

    # vec is a 1-D float32 numpy array (the tensor's storage); off is an element offset
    @staticmethod
    def _p(vec: np.ndarray, off: int) -> int:
        assert vec.dtype == np.float32 and vec.ndim == 1
        return vec.ctypes.data + 4 * int(off)

    def arith(self, name, n, x, xoff, incx, y, yoff, incy):
        getattr(self.lib, "waffe2_s" + name)(int(n), self._p(x, xoff), int(incx), self._p(y, yoff), int(incy))

    def copy(self, n, x, xoff, incx, y, yoff, incy):
        """x <- y.  The reference uses BLAS scopy_ for floats
        (source/backends/cpu/arithmetic.lisp:39-61, 125-142); no arithmetic is involved, so a
        strided numpy assignment has the same result (for incx == 0 the last element wins, as in
        a sequential loop)."""
        n, xoff, incx, yoff, incy = int(n), int(xoff), int(incx), int(yoff), int(incy)
        if n <= 0:
            return
        if incy == 0:
            src = y[yoff]
        elif incy > 0:
            src = y[yoff:yoff + (n - 1) * incy + 1:incy]
        else:
            src = y[yoff + incy * np.arange(n)]
        if incx == 0:
            x[xoff] = src if incy == 0 else src[-1]
        elif incx > 0:
            x[xoff:xoff + (n - 1) * incx + 1:incx] = src
        else:
            x[xoff + incx * np.arange(n)] = src

    def unary(self, name, n, x, xoff, incx, y, yoff, incy):
        getattr(self.lib, "waffe2_s" + name)(int(n), self._p(x, xoff), int(incx), self._p(y, yoff), int(incy))

    def pow(self, n, x, xoff, incx, pow_n, y, yoff, incy):
        self.lib.waffe2_spow(int(n), self._p(x, xoff), int(incx), float(pow_n), self._p(y, yoff), int(incy))

    def maxmin(self, name, n, x, xoff, incx, y, yoff):
        getattr(self.lib, "waffe2_s" + name)(int(n), self._p(x, xoff), int(incx), self._p(y, yoff))

    def cmp(self, name, n, x, xoff, incx, y, yoff, incy, out, ooff, inco, then_v, else_v):
        getattr(self.lib, "waffe2_s" + name)(int(n), self._p(x, xoff), int(incx), self._p(y, yoff), int(incy),
                                             self._p(out, ooff), int(inco), float(then_v), float(else_v))

    def cmp_scal(self, name, n, x, xoff, incx, out, ooff, inco, scal, then_v, else_v):
        getattr(self.lib, "waffe2_s" + name + "_scal")(int(n), self._p(x, xoff), int(incx), self._p(out, ooff),
                                                        int(inco), float(scal), float(then_v), float(else_v))


    def unfold(self, name, col, N, C, H, W, Ho, Wo, KH, KW, pad_h, pad_w, sh, sw, dh, dw, im):
        """`waffe2_{im2col,col2im}_s_column` (cl-waffe2-simd/kernels/im2col.c:63-115,395-448; bound in
        cl-waffe2-simd/api.lisp:27-61,298-310): col is (N C KH KW Ho Wo), im the already padded (N C H W)
        image, both whole contiguous storages."""
        getattr(self.lib, f"waffe2_{name}_s_column")(self._p(col, 0), N, C, H, W, Ho, Wo, KH, KW, pad_h, pad_w, sh, sw, dh, dw,
                                                      self._p(im, 0))


_cache: dict = {}


def load(prefer: str = "reference") -> Kernels:
    """Return the reference build when present (and preferred), else the port."""
    if prefer == "reference" and os.path.exists(REF_SO):
        key = "reference"
    else:
        key = "port"
    if key in _cache:
        return _cache[key]
    if key == "reference":
        k = Kernels(REF_SO, "reference")
    else:
        if not os.path.exists(PORT_SO):
            build()
        k = Kernels(PORT_SO, "port")
    _cache[key] = k
    return k


def have_reference() -> bool:
    return os.path.exists(REF_SO)
