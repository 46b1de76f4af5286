"""`:backward` definitions of the hot-path nodes, as the reference composes them from forward nodes.

In cl-waffe2 a backend supplies FORWARD kernels only: the `:backward` of a `defnode` is itself a small graph of
`!mul` / `!div` / `!cos` ... calls (`/root/reference/source/vm/nodes/defnode.lisp:59-69`), which the compiler lowers to the
same device.  With `(with-devices (CUDATensor) ...)` every step below is therefore one libwaffe2cuda.so kernel.  This
module states those compositions once, over a small backend protocol, so that the device path (`DeviceBackend`, built
on `ops.py`) and the test oracle (an adapter inside `tests/`) run literally the same sequence of node calls:

* AddNode / SubNode / MulNode / DivNode  — `source/base-impl/arithmetic.lisp:83-103`
* unary math nodes                       — `source/base-impl/mathematics.lisp:121-257` (quirks kept as written:
  `cosh' = -sinh`, `sqrt' = 1/x`, `square' = x`, `sign' = x*0`)
* ExptNode                               — `source/base-impl/mathematics.lisp:266-274`
* scalar (op) matrix dispatch            — `source/base-impl/arithmetic.lisp:398-434`
  (`(!sub 1 x)` = `(!add 1 (!scalar-mul -1 x))`, `(!div 1 x)` = `(!mul 1 (!reciprocal x))`,
  `!reciprocal` = DivNode(ones, x), `:179-208`)

Backend protocol (every method returns a fresh tensor unless it ends in `_`):
    copy(x)                      `!copy`
    ones_like(x)                 `(->contiguous (!view (make-tensor 1) (broadcast-to x)))`
    arith_(op, x, y)             X <- X op Y   (y a tensor of x's shape, or a Python number = the scalar-mat nodes)
    unary(fn, x)                 OUT <- f(X)
    expt(x, n)                   OUT <- X^n
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

UNARY_WITH_BACKWARD = ("abs", "sign", "sqrt", "square", "sin", "cos", "tan", "sinh", "cosh", "tanh", "asin", "acos", "atan",
                       "exp", "log2", "log10", "loge", "log1p")


class DeviceBackend:
    """The protocol on CUDATensor: each method is one (or, for `ones_like`, two) C-ABI calls through `ops.py`."""

    def __init__(self):
        from . import ops, tensor
        self._ops, self._tensor = ops, tensor

    def copy(self, x):
        return self._ops.copy(x)

    def ones_like(self, x):
        return self._ops.fill_(self._tensor.make_tensor(x.shape), 1.0)

    def arith_(self, op, x, y):
        return getattr(self._ops, op + "_")(x, y)

    def unary(self, fn, x):
        return self._ops.unary(fn, x)

    def expt(self, x, n):
        return self._ops.expt(x, n)


# ---- the graph builders the :backward bodies call (out of place, like `!mul` etc.)
def _binary(B, op, x, y):
    """`(!op x y)` for tensor/tensor, tensor/number and number/tensor (arithmetic.lisp:398-434)."""
    x_num, y_num = isinstance(x, (int, float)), isinstance(y, (int, float))
    assert not (x_num and y_num)
    if x_num:
        if op == "add" or op == "mul":          # (!add 1 x) -> (!scalar-add 1 x): commutative, scalar second
            return B.arith_(op, B.copy(y), float(x))
        if op == "sub":                         # (!sub 1 x) -> (!add 1 (!scalar-mul -1 x))
            return B.arith_("add", B.arith_("mul", B.copy(y), -1.0), float(x))
        return B.arith_("mul", reciprocal(B, y), float(x))      # (!div 1 x) -> (!mul 1 (!reciprocal x))
    return B.arith_(op, B.copy(x), float(y) if y_num else y)


def reciprocal(B, x):
    """`!reciprocal` (arithmetic.lisp:179-208): DivNode on a dense tensor of ones."""
    return B.arith_("div", B.ones_like(x), x)


def mul(B, x, y): return _binary(B, "mul", x, y)
def div(B, x, y): return _binary(B, "div", x, y)
def add(B, x, y): return _binary(B, "add", x, y)
def sub(B, x, y): return _binary(B, "sub", x, y)


# ---- node backward definitions
def arith_backward(B, op: str, dout, dx, dy) -> Tuple[object, object]:
    """=> (grad of X, grad of Y) for `X <- X op Y` (arithmetic.lisp:83-103)."""
    if op == "add":
        return dout, dout
    if op == "sub":
        return dout, mul(B, -1, dout)
    if op == "mul":
        return mul(B, dout, dy), mul(B, dout, dx)
    if op == "div":
        return div(B, dout, dy), div(B, mul(B, mul(B, dout, -1), dx), mul(B, dy, dy))
    raise ValueError(op)


def unary_backward(B, fn: str, dout, x):
    """=> grad of X for `OUT <- fn(X)` (mathematics.lisp:121-257), the reference's formulas verbatim."""
    u = lambda name, t: B.unary(name, t)
    if fn == "abs":
        return mul(B, dout, u("sign", x))
    if fn == "sign":
        return mul(B, x, 0)                                       # ignores dout, as written
    if fn == "sqrt":
        return mul(B, dout, div(B, 1, x))                         # as written (not 1 / (2 sqrt x))
    if fn == "square":
        return mul(B, dout, x)                                    # as written (not 2x)
    if fn == "sin":
        return mul(B, dout, u("cos", x))
    if fn == "cos":
        return mul(B, dout, mul(B, -1, u("sin", x)))
    if fn == "tan":
        return mul(B, dout, div(B, 1, mul(B, u("cos", x), u("cos", x))))
    if fn == "sinh":
        return mul(B, dout, u("cosh", x))
    if fn == "cosh":
        return mul(B, dout, mul(B, -1, u("sinh", x)))             # as written (sign quirk)
    if fn == "tanh":
        return mul(B, dout, div(B, 1, mul(B, u("cosh", x), u("cosh", x))))
    if fn == "asin":
        return mul(B, dout, div(B, 1, u("sqrt", sub(B, 1, u("square", x)))))
    if fn == "acos":
        return mul(B, dout, div(B, -1, u("sqrt", sub(B, 1, u("square", x)))))
    if fn == "atan":
        return mul(B, dout, div(B, 1, add(B, 1, u("square", x))))
    if fn == "exp":
        return mul(B, dout, u("exp", x))
    if fn == "log2":
        return mul(B, dout, div(B, 1, mul(B, x, math.log(2.0))))
    if fn == "log10":
        return mul(B, dout, div(B, 1, mul(B, x, math.log(10.0))))
    if fn == "loge":
        return mul(B, dout, div(B, 1, x))
    if fn == "log1p":
        return mul(B, dout, div(B, 1, add(B, 1, x)))
    raise ValueError(f"{fn}: no :backward in the reference")


def expt_backward(B, dout, x, n: float, want_dn: bool = False) -> Tuple[object, Optional[object]]:
    """ExptNode (mathematics.lisp:266-274): dX = dout * (n * x^(n-1)); dN (per element, before the reference's
    reduction to the scalar) = dout * (x^n * loge x)."""
    gx = mul(B, dout, mul(B, B.expt(x, n - 1.0), float(n)))
    gn = mul(B, dout, mul(B, B.expt(x, n), B.unary("loge", x))) if want_dn else None
    return gx, gn
