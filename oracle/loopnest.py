"""oracle/loopnest.py — TEST INFRASTRUCTURE, NOT PRODUCT.

CPU restatement of the Lisp side of cl-waffe2's element-wise hot path: the AbstractTensor view
algebra, `solve-loop-order` and the `do-compiled-loop*` nest, driving the reference's own C leaf
kernels (oracle/kernels.py) exactly as the Lisp would: same collapse rule, same per-row calls,
same (n, incx, incy).  Every function cites the reference file:line it follows (paths relative
to /root/reference/source unless noted).

The Lisp layer itself cannot run here (no SBCL in the image), so this file is pinned against the
golden vectors of the reference's own test-suite in tests/test_oracle_golden.py
(`vm/iterator/test-suites.lisp:14-166`, `vm/generic-tensor/t/view.lisp:30-70`,
`base-impl/t/arithmetic.lisp`, `base-impl/t/reduction.lisp`, ...).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import kernels as _k


# ======================================================================================
# Range algebra — vm/iterator/range.lisp
# ======================================================================================
@dataclass
class Range:
    """`(defstruct Range from to step)` — vm/iterator/range.lisp:23-27."""
    frm: int
    to: int
    step: int = 1


def make_range(frm: int, to: Optional[int] = None, step: int = 1) -> Range:
    """`(range from &optional (to (1+ from)) (step 1))` — range.lisp:110-143."""
    if to is None:
        to = frm + 1
    assert step != 0, "range: do not create a range whose step is 0"
    return Range(int(frm), int(to), int(step))


def range_size(r: Range) -> int:
    """range.lisp:46-64: floor(|from - to| / |step|), with the step==1 short cuts."""
    if r.step == 1:
        if r.frm == 0:
            return r.to
        return abs(r.frm - r.to)
    return abs(r.frm - r.to) // abs(r.step)


def range_nth(r: Range, count: int) -> int:
    """range.lisp:72-96.  [a,b) normalised so a<=b (negating the step when swapped); a negative
    step starts from b+step and walks down."""
    a, b, c = r.frm, r.to, r.step
    if a > b:
        a, b = b, a
        c = -c
    if c < 0:
        tmp = b
        b = a
        a = c + tmp
    return a + c * count


def _lazy_min(a, b):
    """range.lisp:12-18 (numeric case)."""
    if a == 0:
        return a
    if b == 0:
        return b
    return min(a, b)


def _lazy_max(a, b):
    """range.lisp:3-10 (numeric case)."""
    if a == 0:
        return b
    if b == 0:
        return a
    return max(a, b)


def _signum(x):
    return (x > 0) - (x < 0)


def compose_range(range2: Range, range1: Optional[Range]) -> Range:
    """`(.range range2 range1)` — range.lisp:158-244: interpret range2 from the viewpoint of
    range1 (range1 = the older view)."""
    if range1 is None:
        return range2
    f1, t1, s1 = range1.frm, range1.to, range1.step
    f2, t2, s2 = range2.frm, range2.to, range2.step
    offset = _lazy_min(f1, t1) if s1 > 0 else _lazy_max(f1, t1)
    frm = offset + _signum(s1) * f2
    to = offset + _signum(s1) * t2
    from1 = _lazy_min(frm, to)
    to1 = _lazy_max(frm, to)
    step = _signum(s2) * math.lcm(s1, s2)
    return make_range(from1, to1, step)


# ======================================================================================
# Subscripts / views — vm/generic-tensor/view.lisp
# ======================================================================================
T = True  # the subscript `t`


def viewtype(view) -> str:
    """view.lisp:95-136 (`with-viewcase` / `viewtype`)."""
    if isinstance(view, (list, tuple)):
        if len(view) == 2 and view[0] in (":broadcast", ":index"):
            return "broadcast" if view[0] == ":broadcast" else "index"
        if len(view) == 2:
            return "slice"
        if len(view) == 3:
            return "slice-step"
        raise ValueError(f"Cannot parse a view {view!r}")
    if isinstance(view, Range):
        return "range"
    if view is True:
        return "t"
    return "index"


def view_to_range(length: int, view) -> Range:
    """view.lisp:138-145."""
    vt = viewtype(view)
    if vt == "index":
        v = view[1] if isinstance(view, (list, tuple)) else view
        return make_range(v)
    if vt in ("slice", "slice-step"):
        return make_range(*view)
    if vt == "t":
        return make_range(0, length)
    if vt == "range":
        return view
    return make_range(0)  # broadcast


@dataclass
class Subscript:
    """view.lisp:71-75."""
    view: object
    broadcast: bool
    range: Range


def _parse_absolute(size, sub):
    """view.lisp:162-166: negative subscripts count from the end of the *viewed* extent."""
    if isinstance(sub, (int, np.integer)) and not isinstance(sub, bool) and sub < 0:
        return size + sub
    return sub


def parse_view(base_size: int, viewed_shape: Optional[int], past_view, view) -> Subscript:
    """view.lisp:168-205.  NOTE (kept as written): the older view enters through its RAW subscript
    (`force-list past-view`), not through its composed range."""
    vt = viewtype(view)
    if vt == "index" and not isinstance(view, (list, tuple)):
        view = _parse_absolute(viewed_shape, view)
    elif vt == "slice":
        view = [_parse_absolute(viewed_shape, view[0]), _parse_absolute(viewed_shape, view[1])]
    elif vt == "slice-step":
        view = [_parse_absolute(viewed_shape, view[0]), _parse_absolute(viewed_shape, view[1]), view[2]]
    latest = view_to_range(viewed_shape if viewed_shape is not None else base_size, view)
    old = view_to_range(base_size, past_view) if past_view is not None else None
    composed = compose_range(latest, old)
    if old is not None and view is True:
        if isinstance(past_view, (list, tuple)) and past_view[0] == ":broadcast":
            return Subscript(past_view, False, old)
        return Subscript(view, False, make_range(0, base_size))
    return Subscript(view, viewtype(view) == "broadcast", composed)


# ======================================================================================
# AbstractTensor restated — vm/generic-tensor/tensor.lisp:40-111
# ======================================================================================
def calc_strides(shape: Sequence[int], order: str = "column") -> List[int]:
    """vm/generic-tensor/utils.lisp:50-67.  `:column` (the default, tensor.lisp:23) is C
    row-major: stride[i] = stride[i+1] * shape[i+1]."""
    n = len(shape)
    st = [1] * n
    if order == "column":
        for i in range(n - 2, -1, -1):
            st[i] = st[i + 1] * shape[i + 1]
    else:
        for i in range(1, n):
            st[i] = st[i - 1] * shape[i - 1]
    return st


class OTensor:
    """The slots of AbstractTensor the hot path reads: vec, orig-shape, stride, view,
    visible-shape, permute-order, initial-offset, order."""

    def __init__(self, vec: np.ndarray, orig_shape, order="column", view=None, past_view=None,
                 create_from: "OTensor" = None):
        assert vec.dtype == np.float32 and vec.ndim == 1
        self.vec = vec
        self.orig_shape = list(orig_shape)
        self.order = order
        self.offset = create_from.offset if create_from is not None else 0
        self.permute_order: List[int] = []
        # initialize-instance :after — tensor.lisp:410-465
        if create_from is not None and create_from.permuted_p() and self.orig_shape == create_from.orig_shape:
            self.stride = list(create_from.stride)
            self.orig_shape = list(create_from.orig_shape)
            self.permute_order = list(create_from.permute_order)
        else:
            self.stride = calc_strides(self.orig_shape, order)
        self.view = self._compute_next_view(past_view, view if view else [T])
        self.visible_shape = self._compute_visible_shape()
        if not self.permute_order:
            n = len(self.visible_shape)
            self.permute_order = list(range(n - 1, -1, -1))

    # view.lisp:147-160
    def _compute_next_view(self, past_view, subscripts) -> List[Subscript]:
        out = []
        for n, base in enumerate(self.orig_shape):
            pv = past_view[n] if past_view is not None and n < len(past_view) else None
            sub = subscripts[n] if n < len(subscripts) else T
            if isinstance(sub, Subscript):
                sub = sub.view
            out.append(parse_view(base,
                                  range_size(pv.range) if pv is not None else None,
                                  pv.view if pv is not None else None,
                                  sub))
        return out

    # view.lisp:214-219
    def _compute_visible_shape(self):
        return [v.view[1] if v.broadcast else range_size(v.range) for v in self.view]

    @property
    def shape(self):
        return list(self.visible_shape)

    def dims(self):
        return len(self.visible_shape)

    def permuted_p(self):
        n = len(self.permute_order)
        return self.permute_order != list(range(n - 1, -1, -1))

    def numel(self):
        return int(np.prod(self.visible_shape)) if self.visible_shape else 1

    # `mref` — tensor.lisp:671-693
    def mref(self, *idx) -> float:
        off = self.offset
        for st, i, v in zip(self.stride, idx, self.view):
            stepby = 0 if viewtype(v.view) == "broadcast" else 1  # compute-stepby, view.lisp:207-212
            off += st * stepby * range_nth(v.range, i)
        return float(self.vec[off])

    def to_numpy(self) -> np.ndarray:
        """Materialise the visible tensor element by element through the mref arithmetic
        (vectorised)."""
        shape = self.visible_shape
        off = np.full(shape if shape else (1,), self.offset, dtype=np.int64)
        for ax, (st, v) in enumerate(zip(self.stride, self.view)):
            if viewtype(v.view) == "broadcast":
                continue
            n = shape[ax]
            idx = np.array([range_nth(v.range, i) for i in range(n)], dtype=np.int64) * st
            sh = [1] * len(shape)
            sh[ax] = n
            off = off + idx.reshape(sh)
        return self.vec[off].reshape(shape)


def make_tensor(shape, data=None, order="column") -> OTensor:
    """`make-tensor` (tensor.lisp:528-610) with LispTensor's zero-filled allocation
    (backends/lisp/tensor.lisp:20-28).  `data` is given in logical C order."""
    n = int(np.prod(shape)) if len(shape) else 1
    t = OTensor(np.zeros(n, dtype=np.float32), shape, order=order)
    if data is not None:
        arr = np.asarray(data, dtype=np.float32).reshape(shape)
        if order == "column":
            t.vec[:] = arr.reshape(-1)
        else:  # :row == Fortran order (utils.lisp:60-67)
            t.vec[:] = arr.reshape(-1, order="F")
    return t


def ax_plus_b(shape, a, b, order="column") -> OTensor:
    """`ax+b` — distributions/sampling.lisp:103-108: element i (flat storage index) = a*i + b."""
    n = int(np.prod(shape))
    t = make_tensor(shape, order=order)
    t.vec[:] = (np.float32(a) * np.arange(n, dtype=np.float32) + np.float32(b)).astype(np.float32)
    return t


def view(t: OTensor, *subscripts) -> OTensor:
    """`view` — tensor.lisp:820-864: new tensor sharing vec; broadcast axes get stride 0."""
    subs = [s.view if isinstance(s, Subscript) else s for s in subscripts]
    out = OTensor(t.vec, t.orig_shape, order=t.order, view=subs, past_view=t.view, create_from=t)
    for n, v in enumerate(out.view):
        if isinstance(v.view, (list, tuple)) and v.view[0] == ":broadcast":
            out.stride[n] = 0
    return out


def permute(t: OTensor, *orders) -> OTensor:
    """`permute*` — tensor.lisp:929-991 with `sync-permute!` (:898-917).  Orders count from the
    back: the initial order is (n-1 ... 1 0); position k takes the axis at `rank - order[k]`."""
    subs = [v.view for v in t.view]
    new = OTensor(t.vec, list(t.orig_shape), order=t.order, view=subs, past_view=list(t.view), create_from=t)
    new.stride = list(t.stride)
    for n, v in enumerate(new.view):
        if isinstance(v.view, (list, tuple)) and v.view[0] == ":broadcast":
            new.stride[n] = 0
    old = t.permute_order
    orders = list(orders)
    if ":~" in orders:
        pos = orders.index(":~")
        pure = [o for o in orders if o != ":~"]
        diff = len(old) - len(pure)
        orders = pure[:pos] + old[pos:pos + diff] + pure[pos:]
    assert sorted(orders) == sorted(old), f"permute*: all axes must be used: {old} -> {orders}"
    new.permute_order = list(orders)
    rank = len(new.visible_shape) - 1

    def apply(lst):
        cp = list(lst)
        return [cp[rank - o] for o in new.permute_order]

    new.stride = apply(new.stride)
    new.view = apply(new.view)
    new.visible_shape = apply(new.visible_shape)
    new.orig_shape = apply(new.orig_shape)
    return new


def transpose(t: OTensor) -> OTensor:
    """`!t` on a matrix == `(permute* x :~ 0 1)` (base-impl/matrix-ops.lisp:97-173 uses a lazy
    flag instead of moving data; as a *view* it swaps the last two axes)."""
    return permute(t, ":~", 0, 1)


# ======================================================================================
# solve-loop-order / do-compiled-loop* — vm/generic-tensor/do-compiled-loop.lisp
# ======================================================================================
@dataclass
class ALoop:
    """`AbstractLoop` — do-compiled-loop.lisp:27-35."""
    rank: int
    by: int
    size: object
    element_n: object
    mode: str  # "batch" | "apply" | "apply-flatten"


def _rest_contiguous_p(rank: int, tensors: Sequence[OTensor]) -> bool:
    """do-compiled-loop.lisp:46-73: the remaining axes can be collapsed iff every remaining view
    is `t` or a broadcast, and per tensor all remaining views have the same view type."""
    for t in tensors:
        views = t.view[rank:]
        for v in views:
            if not (v.view is True or v.broadcast):
                return False
    for t in tensors:
        views = t.view[rank:]
        rep = viewtype(views[0].view)
        if not all(viewtype(v.view) == rep for v in views):
            return False
    return True


def solve_loop_order(tensors: Sequence[OTensor], kernel_size: int = 1, force_order: bool = False) -> List[ALoop]:
    """do-compiled-loop.lisp:76-164 for the element-wise case."""
    assert all(t.dims() >= kernel_size for t in tensors)
    ks = kernel_size - 1
    rank_n = tensors[0].dims()

    def find_size(r):
        return tensors[0].shape[r]

    def compute(rank):
        if (not force_order and ks == 0
                and all(t.permute_order == tensors[0].permute_order for t in tensors)
                and _rest_contiguous_p(rank, tensors)):
            n = 1
            for r in range(rank, rank_n):
                n *= find_size(r)
            return [ALoop(rank=rank, by=1, element_n=n, size=n, mode="apply-flatten")]
        if rank_n - rank == ks + 1:
            n = 1
            for r in range(rank, rank_n):
                n *= find_size(r)
            return [ALoop(rank=rank, by=1, element_n=n, size=1, mode="apply")]
        return [ALoop(rank=rank, by=1, element_n=None, size=find_size(rank), mode="batch")] + compute(rank + 1)

    return compute(0)


@dataclass
class ViewInstruction:
    """view.lisp:50-60: what a leaf kernel receives per tensor: (offset, size, by)."""
    offset: int
    size: int
    by: int


def do_compiled_loop(tensors: Sequence[OTensor], fn: Callable, kernel_size: int = 1,
                     force_order: bool = False, stats: Optional[dict] = None) -> None:
    """do-compiled-loop.lisp:170-298 (run-time twin of call-with-view.lisp:24-229).
    `fn(*per_tensor_views)` is called once per leaf with one ViewInstruction per tensor."""
    plan = solve_loop_order(tensors, kernel_size, force_order)
    offsets0 = [t.offset for t in tensors]

    def helper(c: int, offsets: List[int]):
        if c >= len(plan):
            return
        sub = plan[c]
        rank = sub.rank
        if sub.mode == "batch":
            offsets = list(offsets)
            # :176-190 — add stride * range_nth(view, 0) once (nothing on broadcast axes)
            for pos, t in enumerate(tensors):
                v = t.view[rank]
                start = 0 if v.broadcast else range_nth(v.range, 0)
                if start != 0:
                    offsets[pos] += start * t.stride[rank]
            diffs = []
            for t in tensors:
                v = t.view[rank]
                diffs.append(t.stride[rank] * (range_nth(v.range, 1) - range_nth(v.range, 0)))
            n = sub.size
            for index in range(n):
                helper(c + 1, offsets)
                if index != n - 1:
                    for pos, t in enumerate(tensors):
                        if not t.view[rank].broadcast:
                            offsets[pos] += diffs[pos]
            return
        # :apply / :apply-flatten leaf — :243-295
        args = []
        for pos, t in enumerate(tensors):
            v = t.view[rank]
            off = offsets[pos] + (0 if v.broadcast else t.stride[rank] * range_nth(v.range, 0))
            if sub.mode == "apply-flatten":
                size = sub.element_n
                by = 0 if v.broadcast else 1
            else:
                size = t.shape[rank]
                by = 0 if v.broadcast else t.stride[rank] * (range_nth(v.range, 1) - range_nth(v.range, 0))
            args.append(ViewInstruction(off, size, by))
        if stats is not None:
            stats["leaf_calls"] = stats.get("leaf_calls", 0) + 1
        fn(*args)

    helper(0, offsets0)


# ======================================================================================
# Nodes on the hot path, as CPUTensor runs them
# ======================================================================================
def _K():
    return _k.load()


def arith_(name: str, x: OTensor, y: OTensor, stats=None) -> OTensor:
    """{Add,Sub,Mul,Div}Node CPUTensor impl: X <- X op Y in place
    (backends/cpu/arithmetic.lisp:32-37, 65-123 -> waffe2_s{add,sub,mul,div})."""
    assert x.shape == y.shape, f"shape mismatch {x.shape} vs {y.shape}"
    K = _K()

    def leaf(xv: ViewInstruction, yv: ViewInstruction):
        K.arith(name, yv.size, x.vec, xv.offset, xv.by, y.vec, yv.offset, yv.by)

    do_compiled_loop([x, y], leaf, stats=stats)
    return x


def move_(x: OTensor, y: OTensor, stats=None) -> OTensor:
    """MoveTensorNode: X <- Y via scopy_ (backends/cpu/arithmetic.lisp:39-61, 125-142)."""
    assert x.shape == y.shape
    K = _K()

    def leaf(xv, yv):
        K.copy(yv.size, x.vec, xv.offset, xv.by, y.vec, yv.offset, yv.by)

    do_compiled_loop([x, y], leaf, stats=stats)
    return x


def copy(x: OTensor) -> OTensor:
    """`!copy` = MoveTensorNode into a fresh contiguous tensor (base-impl/fundamental.lisp:80-163)."""
    out = make_tensor(x.shape, order=x.order)
    return move_(out, x)


def binary(name: str, x: OTensor, y: OTensor) -> OTensor:
    """`!add` etc. out of place = `(forward (AddNode) (!copy x) y)`
    (base-impl/arithmetic.lisp:150-152)."""
    return arith_(name, copy(x), y)


def unary(name: str, x: OTensor, out: Optional[OTensor] = None, stats=None) -> OTensor:
    """Unary math node: OUT <- f(X) (backends/cpu/mathematics.lisp:8-53 -> waffe2_s<f>)."""
    if out is None:
        out = make_tensor(x.shape, order=x.order)
    K = _K()

    def leaf(xv, ov):
        K.unary(name, xv.size, x.vec, xv.offset, xv.by, out.vec, ov.offset, ov.by)

    do_compiled_loop([x, out], leaf, stats=stats)
    return out


def expt(x: OTensor, n: float, out: Optional[OTensor] = None) -> OTensor:
    """ExptNode (backends/cpu/mathematics.lisp:55-75 -> waffe2_spow)."""
    if out is None:
        out = make_tensor(x.shape, order=x.order)
    K = _K()

    def leaf(xv, ov):
        K.pow(xv.size, x.vec, xv.offset, xv.by, n, out.vec, ov.offset, ov.by)

    do_compiled_loop([x, out], leaf)
    return out


def where_scal(cmp: str, x: OTensor, scal: float, then_v=1.0, else_v=0.0, out=None) -> OTensor:
    """Where-Operation-Node (backends/cpu/logical.lisp:37-60 -> waffe2_s<cmp>_scal)."""
    if out is None:
        out = make_tensor(x.shape, order=x.order)
    K = _K()

    def leaf(xv, ov):
        K.cmp_scal(cmp, xv.size, x.vec, xv.offset, xv.by, out.vec, ov.offset, ov.by, scal, then_v, else_v)

    do_compiled_loop([x, out], leaf)
    return out


def compare(cmp: str, x: OTensor, y: OTensor, then_v=1.0, else_v=0.0, out=None) -> OTensor:
    """Compare-Operation-Node (backends/cpu/logical.lisp:62-84 -> waffe2_s<cmp>)."""
    if out is None:
        out = make_tensor(x.shape, order=x.order)
    K = _K()

    def leaf(xv, yv, ov):
        K.cmp(cmp, xv.size, x.vec, xv.offset, xv.by, y.vec, yv.offset, yv.by, out.vec, ov.offset, ov.by,
              then_v, else_v)

    do_compiled_loop([x, y, out], leaf)
    return out


def scalar_op_(name: str, x: OTensor, scalar: float) -> OTensor:
    """`A*=scal` etc.: the scalar becomes a 1-element tensor viewed with (:broadcast n) on every
    axis (base-impl/arithmetic.lisp:216-262, fundamental.lisp:525-533)."""
    s = make_tensor([1] * x.dims(), data=[scalar])
    sb = view(s, *[[":broadcast", n] for n in x.shape])
    return arith_(name, x, sb)


def _norm_axes(axis, dims):
    if axis is True or axis is None:
        return list(range(dims))
    if isinstance(axis, int):
        return [axis + dims if axis < 0 else axis]
    return [a + dims if a < 0 else a for a in axis]


def sum_(x: OTensor, axis=True, keepdims=False, div_by: Optional[float] = None, stats=None) -> OTensor:
    """`!sum` — base-impl/reduction.lisp:5-79: out <- out*0 ; out*[:broadcast] += x.
    Returns the reduced tensor with extent-1 axes (keepdims=False) or its broadcast view."""
    dims = x.dims()
    axes = _norm_axes(axis, dims)
    shape = list(x.shape)
    view_args = [T] * dims
    for a in axes:
        view_args[a] = [":broadcast", shape[a]]
        shape[a] = 1
    out = make_tensor(shape, order=x.order)
    scalar_op_("mul", out, 0.0)                 # A*=scal out 0   (:60-63)
    out_b = view(out, *view_args)               # (:71)
    arith_("add", out_b, x, stats=stats)        # A+=B out* x     (:72-79)
    if div_by is not None:
        scalar_op_("div", out, float(div_by))   # !div by the 1-element count tensor (:74-78, 81-136)
    return out_b if keepdims else out


def mean_(x: OTensor, axis=True, keepdims=False) -> OTensor:
    """`!mean` — reduction.lisp:81-136: !sum then divide by the element count (an fp32 value)."""
    axes = _norm_axes(axis, x.dims())
    cnt = np.float32(1.0)
    for a in axes:
        cnt = np.float32(cnt * np.float32(x.shape[a]))
    return sum_(x, axis=axis, keepdims=keepdims, div_by=float(cnt))


def softmax(x: OTensor, avoid_overflow: bool = True, axis=1) -> OTensor:
    """`!softmax` — nn/activation.lisp:126-159 (mean-shift, not max-shift)."""
    if avoid_overflow:
        x1 = binary("sub", x, mean_(x, axis=axis, keepdims=True))
        e = unary("exp", x1)
        z = sum_(e, axis=axis, keepdims=True)
        return arith_("div", copy(e), z)
    e = unary("exp", x)
    z = sum_(e, axis=axis, keepdims=True)
    return arith_("div", copy(e), z)


def maxmin(name: str, x: OTensor) -> OTensor:
    """MaxValue/MinValue over the last axis, one call per row with force-order
    (backends/cpu/matrix-ops.lisp:332-369 -> waffe2_s{max,min})."""
    shape = x.shape[:-1] + [1]
    out = make_tensor(shape, order=x.order)
    ob = view(out, *([T] * (x.dims() - 1) + [[":broadcast", x.shape[-1]]]))
    K = _K()

    def leaf(xv, ov):
        K.maxmin(name, xv.size, x.vec, xv.offset, xv.by, out.vec, ov.offset)

    do_compiled_loop([x, ob], leaf, force_order=True)
    return out


def argmaxmin(name: str, x: OTensor) -> np.ndarray:
    """ArgMax/ArgMin over the last axis via isamax_/isamin_ - 1
    (backends/cpu/matrix-ops.lisp:240-290).  BLAS i?amax ranks |x|; first occurrence wins."""
    arr = x.to_numpy().reshape(-1, x.shape[-1])
    a = np.abs(arr)
    idx = a.argmax(axis=1) if name == "max" else a.argmin(axis=1)
    return idx.astype(np.uint32).reshape(x.shape[:-1] + [1])


# ======================================================================================
# MatMulNode — backends/cpu/matrix-ops.lisp:12-237
# ======================================================================================
def _sgemm(a2: np.ndarray, b2: np.ndarray, ta: bool, tb: bool) -> np.ndarray:
    """One `sgemm_` call exactly as CPUTensor issues it for row-major storage
    (matrix-ops.lisp:58-79): the operands are swapped and the row-major buffers are handed to the
    column-major BLAS untouched, `sgemm_(transB, transA, n, m, k, 1.0, B, ldb, A, lda, 0.0, C, ldc)`.
    OpenBLAS is not vendored by the reference (system `libblas`, no pinned version); the stand-in is
    the OpenBLAS bundled with scipy."""
    from scipy.linalg import blas
    a2 = np.ascontiguousarray(a2, dtype=np.float32)
    b2 = np.ascontiguousarray(b2, dtype=np.float32)
    c_t = blas.sgemm(1.0, b2.T, a2.T, trans_a=1 if tb else 0, trans_b=1 if ta else 0)  # (n, m) column-major
    return np.ascontiguousarray(c_t.T)


def matmul(a: OTensor, b: OTensor, ta: bool = False, tb: bool = False) -> OTensor:
    """`(!matmul a b)` with `(!t ·)` flags: C[~ i k] <- A[~ i j] B[~ j k], one sgemm per batch index
    (`call-with-view :at-least-dim 2`); viewed/permuted operands are made contiguous first."""
    A = a.to_numpy()
    B = b.to_numpy()
    batch_shape = list(A.shape[:-2]) if A.ndim >= B.ndim else list(B.shape[:-2])
    A3 = A.reshape((-1,) + A.shape[-2:])
    B3 = B.reshape((-1,) + B.shape[-2:])
    nb = max(A3.shape[0], B3.shape[0])
    outs = []
    for i in range(nb):
        outs.append(_sgemm(A3[i % A3.shape[0]], B3[i % B3.shape[0]], ta, tb))
    C = np.stack(outs).reshape(batch_shape + list(outs[0].shape))
    return make_tensor(list(C.shape), data=C)
