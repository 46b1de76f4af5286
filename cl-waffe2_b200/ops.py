"""Forward implementations of the hot-path nodes for CUDATensor — what
`source/backends/cuda/{arithmetic,mathematics,logical,matrix-ops,fusion}.lisp` register through
`define-impl-op` (cl-waffe2_b200/lisp/), each one a single call into libwaffe2cuda.so.

Naming follows the reference's `!fn` graph builders with the `!` dropped
(`/root/reference/source/base-impl/arithmetic.lisp:367-434`, `mathematics.lisp:13-257`,
`reduction.lisp:5-136`, `matrix-ops.lisp:6-173`, `nn/activation.lisp:11-159`); `x_` suffix =
the in-place `A+=B` family.  Everything is enqueued asynchronously on the library stream.
"""
from __future__ import annotations

import ctypes as C
import struct
from typing import Optional, Sequence, Union

import numpy as np

from . import _lib
from .tensor import CUDATensor, T, make_tensor, permute, reshape_contiguous, view

Number = Union[int, float]
_OPS = {"add": _lib.ADD, "sub": _lib.SUB, "mul": _lib.MUL, "div": _lib.DIV, "copy": _lib.COPY}


def _L():
    return _lib.load()


def _ref(v):
    return C.byref(v)


def _same_shape(who, a: CUDATensor, b: CUDATensor):
    if a.shape != b.shape:
        # the reference raises a shaping-error from the node's :where clause (vm/nodes/shape-error.lisp)
        raise ValueError(f"{who}: shapes do not match: {a.shape} vs {b.shape} "
                         "(broadcast explicitly with (view x ... (:broadcast n)))")


# ------------------------------------------------------------------ Move / copy
def move_(x: CUDATensor, y: CUDATensor) -> CUDATensor:
    """MoveTensorNode: X <- Y over views (fundamental.lisp:9-43)."""
    _same_shape("MoveTensorNode", x, y)
    _lib.check(_L().wf2cu_ew_binary(_lib.COPY, _ref(x.lower()), _ref(y.lower())))
    return x


def copy(x: CUDATensor) -> CUDATensor:
    """`!copy`: fresh contiguous tensor <- x."""
    out = make_tensor(x.shape, order="column")
    return move_(out, x)


def contiguous(x: CUDATensor) -> CUDATensor:
    """`->contiguous` (fundamental.lisp:1059-1107): copy only when the view is not already dense."""
    return x if (x.is_contiguous() and not x.permuted_p()) else copy(x)


# ------------------------------------------------------------------ arithmetic
def _inplace(op: str, x: CUDATensor, y) -> CUDATensor:
    if isinstance(y, (int, float, np.floating, np.integer)):
        _lib.check(_L().wf2cu_ew_scalar(_OPS[op], _ref(x.lower()), float(y)))
        return x
    _same_shape(f"{op.capitalize()}Node", x, y)
    _lib.check(_L().wf2cu_ew_binary(_OPS[op], _ref(x.lower()), _ref(y.lower())))
    return x


def add_(x, y): return _inplace("add", x, y)      # A+=B  / A+=scal
def sub_(x, y): return _inplace("sub", x, y)      # A-=B
def mul_(x, y): return _inplace("mul", x, y)      # A*=B
def div_(x, y): return _inplace("div", x, y)      # A/=B


def _outofplace(op: str, x, y, out: Optional[CUDATensor] = None) -> CUDATensor:
    """`(!add x y)` = `(forward (AddNode) (!copy x) y)` (arithmetic.lisp:150-152); the copy and
    the in-place op are issued as ONE three-address kernel."""
    if isinstance(x, (int, float)):
        # scalar (op) tensor: (!add 1 x) etc. (arithmetic.lisp:407-426)
        if op in ("add", "mul"):
            return _outofplace(op, y, x, out)
        s = make_tensor([1] * y.dims(), data=[float(x)])
        x = view(s, *[[":broadcast", n] for n in y.shape])
    if isinstance(y, (int, float, np.floating, np.integer)):
        o = out if out is not None else make_tensor(x.shape)
        if o is not x:
            move_(o, x)
        _lib.check(_L().wf2cu_ew_scalar(_OPS[op], _ref(o.lower()), float(y)))
        return o
    _same_shape(f"!{op}", x, y)
    o = out if out is not None else make_tensor(x.shape)
    _lib.check(_L().wf2cu_ew_binary3(_OPS[op], _ref(o.lower()), _ref(x.lower()), _ref(y.lower())))
    return o


def add(x, y, out=None): return _outofplace("add", x, y, out)
def sub(x, y, out=None): return _outofplace("sub", x, y, out)
def mul(x, y, out=None): return _outofplace("mul", x, y, out)
def div(x, y, out=None): return _outofplace("div", x, y, out)


def fill_(x: CUDATensor, value: Number) -> CUDATensor:
    _lib.check(_L().wf2cu_fill(_ref(x.lower()), float(value)))
    return x


# ------------------------------------------------------------------ unary math
def unary(fn: str, x: CUDATensor, out: Optional[CUDATensor] = None) -> CUDATensor:
    """OUT <- f(X) (mathematics.lisp:13-257); `out` defaults to a fresh tensor like `(!sin x)`."""
    if fn not in _lib.UNARY:
        raise ValueError(f"unknown unary node {fn!r}")
    o = out if out is not None else make_tensor(x.shape)
    _same_shape(fn, x, o)
    _lib.check(_L().wf2cu_ew_unary(_lib.UNARY[fn], _ref(x.lower()), _ref(o.lower())))
    return o


def _mk_unary(name):
    def f(x, out=None):
        return unary(name, x, out)
    f.__name__ = name
    f.__doc__ = f"`(!{name} x)`"
    return f


sin, cos, tan = _mk_unary("sin"), _mk_unary("cos"), _mk_unary("tan")
asin, acos, atan = _mk_unary("asin"), _mk_unary("acos"), _mk_unary("atan")
sinh, cosh, tanh = _mk_unary("sinh"), _mk_unary("cosh"), _mk_unary("tanh")
exp, loge, log2, log10, log1p = (_mk_unary(n) for n in ("exp", "loge", "log2", "log10", "log1p"))
sqrt, abs_, sign, square = _mk_unary("sqrt"), _mk_unary("abs"), _mk_unary("sign"), _mk_unary("square")


def expt(x: CUDATensor, n: Number, out=None) -> CUDATensor:
    """ExptNode: OUT <- X^n (mathematics.lisp:267-303)."""
    o = out if out is not None else make_tensor(x.shape)
    _lib.check(_L().wf2cu_pow(_ref(x.lower()), _ref(o.lower()), float(n)))
    return o


# ------------------------------------------------------------------ logical
def where(x: CUDATensor, cmp: str, scal: Number, true_then: Number = 1.0, false_then: Number = 0.0, out=None):
    """Where-Operation-Node, e.g. `(A>scal x 0)` (base-impl/logical.lisp:11-53)."""
    o = out if out is not None else make_tensor(x.shape)
    _lib.check(_L().wf2cu_cmp(_lib.CMP[cmp], _ref(x.lower()), None, float(scal), _ref(o.lower()),
                              float(true_then), float(false_then)))
    return o


def compare(x: CUDATensor, cmp: str, y: CUDATensor, true_then: Number = 1.0, false_then: Number = 0.0, out=None):
    """Compare-Operation-Node, e.g. `(A>B x y)`."""
    _same_shape("Compare-Operation-Node", x, y)
    o = out if out is not None else make_tensor(x.shape)
    _lib.check(_L().wf2cu_cmp(_lib.CMP[cmp], _ref(x.lower()), _ref(y.lower()), 0.0, _ref(o.lower()),
                              float(true_then), float(false_then)))
    return o


# ------------------------------------------------------------------ reductions
def _axes(axis, dims):
    if axis is True or axis is None:
        return list(range(dims))
    if isinstance(axis, (int, np.integer)):
        return [int(axis) + dims if axis < 0 else int(axis)]
    return [int(a) + dims if a < 0 else int(a) for a in axis]


def sum(x: CUDATensor, axis=True, keepdims: bool = False, out: Optional[CUDATensor] = None,
        _divisor: float = 0.0) -> CUDATensor:
    """`!sum` (reduction.lisp:5-79).  The reference's three instructions (zero-fill by `A*=scal`,
    broadcast view, `A+=B`) are one fused reduction kernel here."""
    dims = x.dims()
    axes = _axes(axis, dims)
    shape = x.shape
    vargs = [T] * dims
    for a in axes:
        vargs[a] = [":broadcast", shape[a]]
        shape[a] = 1
    o = out if out is not None else make_tensor(shape)
    if o.shape != shape:
        raise ValueError(f"!sum: the given out's shape is {o.shape}, but expected: {shape}")
    ob = view(o, *vargs)
    _lib.check(_L().wf2cu_reduce_sum(_ref(x.lower()), _ref(ob.lower()), 0, float(_divisor)))
    return ob if keepdims else o


def mean(x: CUDATensor, axis=True, keepdims: bool = False, out=None) -> CUDATensor:
    """`!mean` (reduction.lisp:81-136): sum then a true divide by the element count."""
    cnt = np.float32(1.0)
    for a in _axes(axis, x.dims()):
        cnt = np.float32(cnt * np.float32(x.shape[a]))
    return sum(x, axis=axis, keepdims=keepdims, out=out, _divisor=float(cnt))


def softmax(x: CUDATensor, avoid_overflow: bool = True, axis: int = 1, out=None, pre_divisor: float = 0.0) -> CUDATensor:
    """`!softmax` (nn/activation.lisp:126-159) as one kernel.  Mean-shift, as in the reference.
    pre_divisor: fold a `(!div x n)` node in front into the same pass (wf2cu_softmax_ex); when the library does
    not implement it for this layout the two nodes are issued separately, exactly as written."""
    o = out if out is not None else make_tensor(x.shape)
    if pre_divisor:
        rc = _L().wf2cu_softmax_ex(_ref(x.lower()), _ref(o.lower()), int(axis), 1 if avoid_overflow else 0, float(pre_divisor))
        if rc == 0:
            return o
        if rc != 3:                              # WF2CU_ERR_UNSUPPORTED -> the unfused sequence
            _lib.check(rc)
        x = div(x, float(pre_divisor))
    _lib.check(_L().wf2cu_softmax(_ref(x.lower()), _ref(o.lower()), int(axis), 1 if avoid_overflow else 0))
    return o


def _rowred(x: CUDATensor, is_min: bool, arg: bool):
    shape = x.shape[:-1] + [1]
    o = make_tensor(shape)
    fn = _L().wf2cu_argmaxmin if arg else _L().wf2cu_reduce_maxmin
    _lib.check(fn(1 if is_min else 0, _ref(x.lower()), _ref(o.lower())))
    return o


def max(x): return _rowred(x, False, False)       # MaxValueNode over the last axis
def min(x): return _rowred(x, True, False)


def argmax(x) -> np.ndarray:
    """ArgMaxNode over the last axis; returns uint32 indices (matrix-ops.lisp:269-271) on the host."""
    o = _rowred(x, False, True)
    return o.to_numpy().view(np.uint32)


def argmin(x) -> np.ndarray:
    o = _rowred(x, True, True)
    return o.to_numpy().view(np.uint32)


# ------------------------------------------------------------------ matmul
class LazyT:
    """`(!t x)`: zero-copy transpose flag consumed by `!matmul` (LazyTransposeNode,
    base-impl/matrix-ops.lisp:97-173)."""
    def __init__(self, x: CUDATensor):
        self.x = x


def t(x):
    return x.x if isinstance(x, LazyT) else LazyT(x)


def _mat_operand(x: CUDATensor):
    """Return (tensor, ld, batch_shape, batch_stride) with the last two axes dense-rowed and the
    batch axes collapsible to one stride; otherwise make the operand contiguous first
    (`adjust-layout`, backends/cpu/matrix-ops.lisp)."""
    for attempt in range(2):
        lv = x.lower()
        r = lv.rank
        ok = r >= 2 and (lv.stride[r - 1] == 1 or lv.shape[r - 1] == 1) and lv.stride[r - 2] >= lv.shape[r - 1]
        bstride = 0
        if ok and r > 2:
            # batch axes must collapse into one (possibly broadcast) stride
            bstride = lv.stride[r - 3]
            run = bstride * lv.shape[r - 3]
            for i in range(r - 4, -1, -1):
                if lv.shape[i] != 1 and lv.stride[i] != run:
                    ok = False
                run = lv.stride[i] * lv.shape[i] if lv.shape[i] != 1 else run
        if ok:
            batch = 1
            for i in range(r - 2):
                batch *= lv.shape[i]
            return x, lv, int(lv.stride[r - 2]), batch, int(bstride)
        x = copy(x)
    raise AssertionError("unreachable")


def matmul(a, b, out: Optional[CUDATensor] = None, bias: Optional[CUDATensor] = None, relu: bool = False,
           relu_mask: Optional[CUDATensor] = None, gelu: bool = False, residual: Optional[CUDATensor] = None) -> CUDATensor:
    """`(!matmul a b)` with `(!t ·)` operands (base-impl/matrix-ops.lisp:6-173):
    C[~ i k] <- A[~ i j] . B[~ j k], alpha=1, beta=0.
    bias / relu / relu_mask fold the node that follows into the epilogue (wf2cu_gemm_ex): `+ bias` per column
    (LinearLayer, nn/regression.lisp:14-18), `!relu` of the result, or `* (mask > 0)` (the `!relu` backward).
    gelu / residual (wf2cu_gemm_ex2): `!gelu` of the result, `(!add residual result)` — the GPT-2 block's two patterns."""
    ta = isinstance(a, LazyT)
    tb = isinstance(b, LazyT)
    A = a.x if ta else a
    B = b.x if tb else b
    if A.dims() < 2 or B.dims() < 2:
        raise ValueError("!matmul: operands must have rank >= 2")
    A, la, lda, batch_a, sa = _mat_operand(A)
    B, lb, ldb, batch_b, sb = _mat_operand(B)
    ar, ac = A.shape[-2], A.shape[-1]
    br, bc = B.shape[-2], B.shape[-1]
    m, k = (ac, ar) if ta else (ar, ac)
    k2, n = (bc, br) if tb else (br, bc)
    if k != k2:
        raise ValueError(f"!matmul: inner dimensions do not match: {A.shape}{'^T' if ta else ''} x {B.shape}{'^T' if tb else ''}")
    if batch_a != batch_b and 1 not in (batch_a, batch_b):
        raise ValueError(f"!matmul: batch dimensions do not match: {A.shape} vs {B.shape}")
    batch = batch_a if batch_a >= batch_b else batch_b
    if batch_a == 1 and batch > 1:
        sa = 0
    if batch_b == 1 and batch > 1:
        sb = 0
    bshape = A.shape[:-2] if batch_a >= batch_b else B.shape[:-2]
    o = out if out is not None else make_tensor(bshape + [m, n])
    if o.shape != bshape + [m, n] or not o.is_contiguous():
        raise ValueError(f"!matmul: out must be a contiguous {bshape + [m, n]} tensor, got {o.shape}")
    lo = o.lower()
    if gelu or residual is not None:
        if batch != 1 or relu_mask is not None or (gelu and relu):
            raise ValueError("!matmul: the gelu / residual epilogue takes unbatched operands and excludes relu / relu_mask")
        bp, boff = (None, 0)
        if bias is not None:
            if bias.numel() != n or not bias.is_contiguous():
                raise ValueError(f"!matmul: bias must be a contiguous {n}-element tensor, got {bias.shape}")
            bp, boff = bias.vec.ptr, bias.lower().offset
        rp, roff, ldr = (None, 0, 0)
        if residual is not None:
            if residual.numel() != m * n or not residual.is_contiguous():
                raise ValueError(f"!matmul: residual must be a contiguous {[m, n]} tensor, got {residual.shape}")
            rp, roff, ldr = residual.vec.ptr, residual.lower().offset, n
        _lib.check(_L().wf2cu_gemm_ex2(1 if ta else 0, 1 if tb else 0, m, n, k, A.vec.ptr, la.offset, lda,
                                       B.vec.ptr, lb.offset, ldb, o.vec.ptr, lo.offset, n, bp, boff, 2 if gelu else (1 if relu else 0),
                                       rp, roff, ldr))
        return o
    if bias is not None or relu or relu_mask is not None:
        if batch != 1:
            raise ValueError("!matmul: the fused epilogue takes unbatched operands")
        bp, boff = (None, 0)
        if bias is not None:
            if bias.numel() != n or not bias.is_contiguous():
                raise ValueError(f"!matmul: bias must be a contiguous {n}-element tensor, got {bias.shape}")
            bp, boff = bias.vec.ptr, bias.lower().offset
        mp, moff, ldm = (None, 0, 0)
        if relu_mask is not None:
            if relu_mask.shape != [m, n] or not relu_mask.is_contiguous():
                raise ValueError(f"!matmul: relu_mask must be a contiguous {[m, n]} tensor, got {relu_mask.shape}")
            mp, moff, ldm = relu_mask.vec.ptr, relu_mask.lower().offset, n
        _lib.check(_L().wf2cu_gemm_ex(1 if ta else 0, 1 if tb else 0, m, n, k, A.vec.ptr, la.offset, lda,
                                      B.vec.ptr, lb.offset, ldb, o.vec.ptr, lo.offset, n, bp, boff, 1 if relu else 0, mp, moff, ldm))
        return o
    _lib.check(_L().wf2cu_gemm(1 if ta else 0, 1 if tb else 0, m, n, k,
                               A.vec.ptr, la.offset, lda, B.vec.ptr, lb.offset, ldb,
                               o.vec.ptr, lo.offset, n, batch, sa, sb, m * n))
    return o


# ------------------------------------------------------------------ composites (nn/activation.lisp)
def _imm(x: float) -> int:
    return struct.unpack("<i", struct.pack("<f", float(x)))[0]


def fused(program: Sequence[Sequence[int]], views: Sequence[CUDATensor]) -> None:
    """Run a fused element-wise register program over equal-shape views (wf2cu_fused)."""
    flat = (C.c_int32 * (4 * len(program)))(*[int(v) for ins in program for v in ins])
    arr = (_lib.View * len(views))(*[v.lower() for v in views])
    _lib.check(_L().wf2cu_fused(flat, len(program), arr, len(views)))


def sigmoid(x: CUDATensor, out=None) -> CUDATensor:
    """`!sigmoid` = 1 / (1 + exp(-x)) (nn/activation.lisp:43), one pass."""
    o = out if out is not None else make_tensor(x.shape)
    P = _lib
    prog = [(P.P_LOAD, 0, 0, 0), (P.P_UNARY, 1, 0, P.UNARY["neg"]), (P.P_UNARY, 1, 1, P.UNARY["exp"]),
            (P.P_IMM, 2, 0, _imm(1.0)), (P.P_ADD, 1, 2, 1), (P.P_DIV, 1, 2, 1), (P.P_STORE, 1, 1, 0)]
    fused(prog, [x, o])
    return o


def relu(x: CUDATensor, out=None) -> CUDATensor:
    """`!relu` = x * (x > 0 ? 1 : 0) (nn/activation.lisp:11-26), one pass."""
    o = out if out is not None else make_tensor(x.shape)
    P = _lib
    prog = [(P.P_LOAD, 0, 0, 0), (P.P_IMM, 1, 0, _imm(0.0)), (P.P_CMPSEL | (P.CMP["gt"] << 8), 2, 0, 1),
            (P.P_MUL, 2, 0, 2), (P.P_STORE, 2, 1, 0)]
    fused(prog, [x, o])
    return o


def gelu(x: CUDATensor, out=None) -> CUDATensor:
    """`!gelu` tanh approximation as written in nn/activation.lisp:61-66:
    0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))."""
    o = out if out is not None else make_tensor(x.shape)
    P = _lib
    c = float(np.float32(np.sqrt(2.0 / np.pi)))
    prog = [(P.P_LOAD, 0, 0, 0),
            (P.P_MUL, 1, 0, 0), (P.P_MUL, 1, 1, 0),                    # x^3
            (P.P_IMM, 2, 0, _imm(0.044715)), (P.P_MUL, 1, 1, 2), (P.P_ADD, 1, 0, 1),
            (P.P_IMM, 2, 0, _imm(c)), (P.P_MUL, 1, 2, 1), (P.P_UNARY, 1, 1, P.UNARY["tanh"]),
            (P.P_IMM, 2, 0, _imm(1.0)), (P.P_ADD, 1, 2, 1),
            (P.P_IMM, 2, 0, _imm(0.5)), (P.P_MUL, 2, 2, 0), (P.P_MUL, 1, 2, 1),
            (P.P_STORE, 1, 1, 0)]
    fused(prog, [x, o])
    return o


def sync() -> None:
    _lib.check(_L().wf2cu_sync())
