"""CUDATensor — host-side mirror of the device class `source/backends/cuda/tensor.lisp` would
define (see cl-waffe2_b200/lisp/), following the AbstractTensor contract of
`/root/reference/source/vm/generic-tensor/tensor.lisp:40-111,146-210`:

* `vec` is a first-class storage handle (refcounted in C: the VM aliases storage with
  `(setf (tensor-vec a) (tensor-vec b))`, `vm/optimize-ir.lisp:4-18`), zero-filled on allocation
  like LispTensor (`backends/lisp/tensor.lisp:20-28`);
* views (`view`, `permute`) share the vec and only change the descriptor;
* `lower()` folds the reference's view algebra (Range/Subscript, `vm/iterator/range.lisp:72-96`,
  `vm/generic-tensor/view.lisp:138-205`) into one `wf2cu_view` = {base, offset, shape[], stride[]}:
  the whole view goes to the kernel instead of one innermost row at a time
  (`do-compiled-loop.lisp:170-298`).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional, Sequence

import numpy as np

from . import _lib

T = True  # the subscript `t`


# --------------------------------------------------------------------------- storage
class Storage:
    """Device buffer handle (the `vec` slot).  Freed when the last Python reference dies
    (the Lisp side uses `tensor-finalizer` + trivial-garbage, tensor.lisp:323-329)."""

    __slots__ = ("ptr", "nbytes", "_owned", "__weakref__")

    def __init__(self, nbytes: int, ptr: Optional[int] = None):
        self.nbytes = int(nbytes)
        if ptr is None:
            out = C.c_void_p()
            _lib.check(_lib.load().wf2cu_malloc(self.nbytes, C.byref(out)))
            self.ptr = out.value
            self._owned = True
        else:  # adopt a caller-owned device pointer (`:vec` initarg)
            self.ptr = int(ptr)
            self._owned = False

    def __del__(self):
        try:
            if self._owned and self.ptr and _lib._lib is not None and _lib._inited_device is not None:
                _lib._lib.wf2cu_release(self.ptr)
        except Exception:
            pass
        self.ptr = None


# --------------------------------------------------------------------------- range algebra
def _rng(frm, to=None, step=1):
    if to is None:
        to = frm + 1
    if step == 0:
        raise ValueError("range: step must not be 0")
    return (int(frm), int(to), int(step))


def range_size(r) -> int:
    f, t, s = r
    if s == 1:
        return t if f == 0 else abs(f - t)
    return abs(f - t) // abs(s)


def range_nth(r, i: int) -> int:
    lo, hi, s = r
    if lo > hi:
        lo, hi, s = hi, lo, -s
    if s < 0:
        lo = hi + s
    return lo + s * i


def _zmin(a, b):
    return a if a == 0 else (b if b == 0 else min(a, b))


def _zmax(a, b):
    return b if a == 0 else (a if b == 0 else max(a, b))


def _compose(new, old):
    if old is None:
        return new
    f1, t1, s1 = old
    f2, t2, s2 = new
    sg1 = (s1 > 0) - (s1 < 0)
    sg2 = (s2 > 0) - (s2 < 0)
    origin = _zmin(f1, t1) if s1 > 0 else _zmax(f1, t1)
    a, b = origin + sg1 * f2, origin + sg1 * t2
    return _rng(_zmin(a, b), _zmax(a, b), sg2 * math.lcm(s1, s2))


def _kind(sub) -> str:
    if isinstance(sub, (list, tuple)):
        if len(sub) == 2 and sub[0] == ":broadcast":
            return "broadcast"
        if len(sub) == 2 and sub[0] == ":index":
            return "index"
        if len(sub) in (2, 3):
            return "slice"
        raise ValueError(f"cannot parse view subscript {sub!r}: expected (start stop [step])")
    if sub is True:
        return "t"
    if isinstance(sub, (int, np.integer)):
        return "index"
    raise ValueError(f"unknown view subscript {sub!r}")


def _sub_range(extent, sub):
    k = _kind(sub)
    if k == "index":
        return _rng(sub[1] if isinstance(sub, (list, tuple)) else sub)
    if k == "slice":
        return _rng(*sub)
    if k == "t":
        return _rng(0, extent)
    return _rng(0)


class Sub:
    """One axis of a view: raw subscript, broadcast flag, composed range (view.lisp:71-75)."""
    __slots__ = ("raw", "bcast", "rng")

    def __init__(self, raw, bcast, rng):
        self.raw, self.bcast, self.rng = raw, bcast, rng


def _absolute(extent, x):
    return extent + x if isinstance(x, (int, np.integer)) and not isinstance(x, bool) and x < 0 else x


def _next_sub(base_extent: int, prev: Optional[Sub], sub) -> Sub:
    viewed = range_size(prev.rng) if prev is not None else None
    k = _kind(sub)
    if k == "index" and not isinstance(sub, (list, tuple)):
        sub = _absolute(viewed, sub)
    elif k == "slice":
        sub = [_absolute(viewed, sub[0]), _absolute(viewed, sub[1])] + list(sub[2:])
    latest = _sub_range(viewed if viewed is not None else base_extent, sub)
    old = _sub_range(base_extent, prev.raw) if prev is not None else None
    if old is not None and sub is True:
        if _kind(prev.raw) == "broadcast":
            return Sub(prev.raw, False, old)
        return Sub(True, False, _rng(0, base_extent))
    return Sub(sub, k == "broadcast", _compose(latest, old))


def calc_strides(shape: Sequence[int], order: str = "column") -> List[int]:
    """`:column` (default) = C row-major; `:row` = Fortran order
    (vm/generic-tensor/utils.lisp:50-67)."""
    n = len(shape)
    st = [1] * n
    if order == "column":
        for i in range(n - 2, -1, -1):
            st[i] = st[i + 1] * shape[i + 1]
    elif order == "row":
        for i in range(1, n):
            st[i] = st[i - 1] * shape[i - 1]
    else:
        raise ValueError(f"unknown order {order!r}")
    return st


# --------------------------------------------------------------------------- the tensor
class CUDATensor:
    """fp32 device tensor with cl-waffe2 view semantics."""

    def __init__(self, vec: Storage, orig_shape, order="column", subs=None, prev=None, source: "CUDATensor" = None,
                 requires_grad=False):
        self.vec = vec
        self.orig_shape = [int(s) for s in orig_shape]
        self.order = order
        self.dtype = "float"
        self.initial_offset = source.initial_offset if source is not None else 0
        self.requires_grad = requires_grad
        self.grad: Optional["CUDATensor"] = None
        if source is not None and source.permuted_p() and self.orig_shape == source.orig_shape:
            self.stride = list(source.stride)
            self.permute_order = list(source.permute_order)
        else:
            self.stride = calc_strides(self.orig_shape, order)
            self.permute_order = list(range(len(self.orig_shape) - 1, -1, -1))
        subs = list(subs) if subs else []
        self.view: List[Sub] = []
        for ax, ext in enumerate(self.orig_shape):
            p = prev[ax] if prev is not None and ax < len(prev) else None
            s = subs[ax] if ax < len(subs) else T
            self.view.append(_next_sub(ext, p, s.raw if isinstance(s, Sub) else s))
        self.visible_shape = [v.raw[1] if v.bcast else range_size(v.rng) for v in self.view]

    # ---- reference accessors
    @property
    def shape(self) -> List[int]:
        return list(self.visible_shape)

    def dims(self) -> int:
        return len(self.visible_shape)

    def numel(self) -> int:
        n = 1
        for s in self.visible_shape:
            n *= s
        return n

    def permuted_p(self) -> bool:
        n = len(self.permute_order)
        return self.permute_order != list(range(n - 1, -1, -1))

    def tensor_vec(self) -> Storage:
        return self.vec

    # ---- lowering
    def lower(self) -> _lib.View:
        """AbstractTensor -> wf2cu_view (include/waffe2cuda.h)."""
        off = self.initial_offset
        shape, stride = [], []
        for st, v, ext in zip(self.stride, self.view, self.visible_shape):
            shape.append(ext)
            if v.bcast or _kind(v.raw) == "broadcast":
                stride.append(0)
            else:
                first = range_nth(v.rng, 0)
                off += st * first
                stride.append(st * (range_nth(v.rng, 1) - first))
        # a kernel trusts the descriptor: refuse one that addresses elements outside the storage (possible through
        # `permute*` of a sliced view, which re-applies the slice: tensor.lisp:880-896, 929-991) instead of faulting
        if all(ext > 0 for ext in shape):
            lo = off + sum((ext - 1) * sd for ext, sd in zip(shape, stride) if sd < 0)
            hi = off + sum((ext - 1) * sd for ext, sd in zip(shape, stride) if sd > 0)
            if lo < 0 or hi * 4 >= self.vec.nbytes:
                raise ValueError(f"view of shape {shape} addresses elements [{lo}, {hi}] outside its storage of "
                                 f"{self.vec.nbytes // 4} elements")
        return _lib.make_view(self.vec.ptr, off, shape, stride)

    def is_contiguous(self) -> bool:
        lv = self.lower()
        exp = 1
        for i in range(lv.rank - 1, -1, -1):
            if lv.shape[i] != 1 and lv.stride[i] != exp:
                return False
            exp *= lv.shape[i]
        return True

    # ---- element access (`vref` / `mref`, tensor.lisp:671-746): single-element D2H, syncs
    def vref(self, index: int) -> float:
        out = C.c_float()
        _lib.check(_lib.load().wf2cu_get_f32(self.vec.ptr, int(index), C.byref(out)))
        return out.value

    def set_vref(self, index: int, value: float) -> None:
        _lib.check(_lib.load().wf2cu_set_f32(self.vec.ptr, int(index), float(value)))

    def mref(self, *idx) -> float:
        lv = self.lower()
        off = lv.offset
        for i, k in enumerate(idx):
            off += lv.stride[i] * k
        return self.vref(off)

    # ---- bulk transfer (`convert-tensor-facet`, array-converter.lisp:154-218)
    def to_numpy(self) -> np.ndarray:
        """Visible tensor as a C-ordered numpy array (one bulk D2H of the storage + host gather)."""
        host = np.empty(self.vec.nbytes // 4, dtype=np.float32)
        _lib.check(_lib.load().wf2cu_d2h(self.vec.ptr, 0, host.ctypes.data, host.nbytes))
        lv = self.lower()
        shape = [lv.shape[i] for i in range(lv.rank)]
        strides = [lv.stride[i] * 4 for i in range(lv.rank)]
        lo = lv.offset + sum(min(0, lv.stride[i] * (lv.shape[i] - 1)) for i in range(lv.rank))
        assert lo >= 0
        view = np.lib.stride_tricks.as_strided(host[lv.offset:], shape=shape, strides=strides) if shape else host[lv.offset:lv.offset + 1]
        return np.array(view, dtype=np.float32, copy=True)

    def copy_from_numpy(self, arr) -> "CUDATensor":
        """Upload `arr` (logical C order) into this tensor's whole storage; tensor must be an
        un-viewed allocation."""
        a = np.asarray(arr, dtype=np.float32).reshape(self.orig_shape)
        flat = a.reshape(-1) if self.order == "column" else a.reshape(-1, order="F")
        flat = np.ascontiguousarray(flat)
        _lib.check(_lib.load().wf2cu_h2d(self.vec.ptr, self.initial_offset * 4, flat.ctypes.data, flat.nbytes))
        _lib.check(_lib.load().wf2cu_sync())  # `flat` may be a temporary
        return self

    def __repr__(self):
        return f"<CUDATensor {self.shape} order={self.order} permute={self.permute_order}>"


# --------------------------------------------------------------------------- static allocation
class StaticPool:
    """Static memory plan of a compiled step — the analogue of `VMAllocation` /
    `with-static-allocation` (`source/vm/allocation.lisp:46-171`): the first run under the pool
    allocates, every later run hands out the same buffers in the same order, so the launch
    sequence has fixed addresses and can be captured into a CUDA graph.  Like the reference's
    pool, reused buffers are NOT re-zeroed; every producer overwrites its output."""

    def __init__(self):
        self.slots = []      # [(nbytes, Storage)]
        self.cursor = 0
        self.frozen = False

    def take(self, nbytes: int) -> Storage:
        if self.cursor < len(self.slots):
            nb, st = self.slots[self.cursor]
            if nb != nbytes:
                raise RuntimeError(f"StaticPool: allocation #{self.cursor} changed size ({nb} -> {nbytes}); "
                                   "the compiled step is not static")
            self.cursor += 1
            return st
        if self.frozen:
            raise RuntimeError("StaticPool: new allocation after the plan was frozen (inside graph capture)")
        st = Storage(nbytes)
        self.slots.append((nbytes, st))
        self.cursor += 1
        return st

    def __enter__(self):
        global _active_pool
        self._prev = _active_pool
        _active_pool = self
        self.cursor = 0
        return self

    def __exit__(self, *exc):
        global _active_pool
        _active_pool = self._prev
        return False

    def nbytes(self) -> int:
        return sum(nb for nb, _ in self.slots)


_active_pool = None


# --------------------------------------------------------------------------- constructors
def make_tensor(shape, data=None, order="column", requires_grad=False, initial_element=None) -> CUDATensor:
    """`make-tensor` (tensor.lisp:528-610): allocate zero-filled device storage."""
    _lib.init(_lib._inited_device if _lib._inited_device is not None else 0)
    shape = [int(s) for s in shape]
    n = 1
    for s in shape:
        n *= s
    nbytes = max(n, 1) * 4
    st = _active_pool.take(nbytes) if (_active_pool is not None and data is None and initial_element is None) \
        else Storage(nbytes)
    t = CUDATensor(st, shape, order=order, requires_grad=requires_grad)
    if data is not None:
        t.copy_from_numpy(data)
    elif initial_element not in (None, 0, 0.0):
        _lib.check(_lib.load().wf2cu_fill(C.byref(t.lower()), float(initial_element)))
    return t


def from_numpy(arr, order="column", requires_grad=False) -> CUDATensor:
    a = np.asarray(arr, dtype=np.float32)
    return make_tensor(a.shape, data=a, order=order, requires_grad=requires_grad)


def ax_plus_b(shape, a, b, order="column") -> CUDATensor:
    """`ax+b` (distributions/sampling.lisp:103-108): storage element i = a*i + b, built on the host
    and uploaded in one copy (SURVEY §7 hard part h)."""
    n = int(np.prod(shape))
    t = make_tensor(shape, order=order)
    host = (np.float32(a) * np.arange(n, dtype=np.float32) + np.float32(b)).astype(np.float32)
    _lib.check(_lib.load().wf2cu_h2d(t.vec.ptr, 0, host.ctypes.data, host.nbytes))
    _lib.check(_lib.load().wf2cu_sync())
    return t


def view(t: CUDATensor, *subscripts) -> CUDATensor:
    """`view` (tensor.lisp:820-864)."""
    subs = [s.raw if isinstance(s, Sub) else s for s in subscripts]
    out = CUDATensor(t.vec, t.orig_shape, order=t.order, subs=subs, prev=t.view, source=t)
    for ax, v in enumerate(out.view):
        if _kind(v.raw) == "broadcast":
            out.stride[ax] = 0
    return out


def permute(t: CUDATensor, *orders) -> CUDATensor:
    """`permute*` (tensor.lisp:929-991).  Orders count from the back: the identity is
    (n-1 ... 1 0); `(permute t ":~", 0, 1)` swaps the last two axes."""
    new = CUDATensor(t.vec, t.orig_shape, order=t.order, subs=[v.raw for v in t.view], prev=t.view, source=t)
    new.stride = [0 if _kind(v.raw) == "broadcast" else s for s, v in zip(t.stride, new.view)]
    old = list(t.permute_order)
    orders = list(orders)
    if orders.count(":~") > 1:
        raise ValueError("permute*: the keyword :~ must appear at most once")
    if ":~" in orders:
        pos = orders.index(":~")
        pure = [o for o in orders if o != ":~"]
        gap = len(old) - len(pure)
        orders = pure[:pos] + old[pos:pos + gap] + pure[pos:]
    if sorted(orders) != sorted(old):
        raise ValueError(f"permute*: before and after the operation, all axes must be used: {old} -> {orders}")
    top = len(old) - 1
    pick = [top - o for o in orders]
    new.permute_order = orders
    new.stride = [new.stride[p] for p in pick]
    new.view = [new.view[p] for p in pick]
    new.visible_shape = [new.visible_shape[p] for p in pick]
    new.orig_shape = [new.orig_shape[p] for p in pick]
    return new


def transpose(t: CUDATensor) -> CUDATensor:
    """`!t` as a view (swap the last two axes)."""
    return permute(t, ":~", 0, 1)


def reshape_contiguous(t: CUDATensor, shape) -> CUDATensor:
    """`!reshape` of a contiguous tensor: same vec, new orig-shape (fundamental.lisp:236-297)."""
    shape = list(shape)
    if -1 in shape or True in shape:
        known = 1
        for s in shape:
            if s not in (-1, True):
                known *= s
        shape = [t.numel() // known if s in (-1, True) else s for s in shape]
    n = 1
    for s in shape:
        n *= s
    if n != t.numel():
        raise ValueError(f"!reshape: total size must not change: {t.shape} -> {shape}")
    if not t.is_contiguous():
        raise ValueError("!reshape: tensor must be contiguous (call ->contiguous first)")
    out = CUDATensor(t.vec, shape, order="column")
    out.initial_offset = t.lower().offset
    return out


def tensor_at(storage: Storage, offset: int, shape, requires_grad=False) -> CUDATensor:
    """A contiguous tensor living at element `offset` of an existing storage (how parameters and
    gradients are laid out in one arena so the optimizer and the allreduce see ONE bucket)."""
    t = CUDATensor(storage, [int(s) for s in shape], order="column", requires_grad=requires_grad)
    t.initial_offset = int(offset)
    return t


def as_strided(t: CUDATensor, shape, strides, offset: int = 0) -> CUDATensor:
    """A tensor over t's storage with explicit element strides — what a chain of `!view` (slice) +
    `!reshape` + `!permute` leaves in `tensor-stride` / `tensor-initial-offset`
    (examples/gpt-2/utils.lisp:204-220 `split-heads`), built in one step.  `offset` is relative to
    t's own base offset; t must be a dense tensor."""
    base = t.lower().offset
    out = CUDATensor(t.vec, [int(s) for s in shape], order="column")
    out.stride = [int(s) for s in strides]
    out.initial_offset = int(base + offset)
    return out
