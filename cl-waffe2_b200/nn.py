"""Training-step layer: reverse-mode nodes, `LinearLayer`, losses, optimizers and the compiled
(CUDA-graph) step for the MLP configs (C1 `examples/mlp_sin_wave.lisp`, C4 `examples/mnist/mlp.lisp`).

This mirrors what runs *unchanged* on top of the backend in the reference — `cl-waffe2/nn`
(`/root/reference/source/nn/regression.lisp:14-55`, `activation.lisp:11-43`, `criterion.lisp:53-188`) and
`cl-waffe2/optimizers` (`optimizers/impls/{sgd,adam}.lisp`) — restated in Python over the C ABI so the
step can be driven and measured here.  Backward formulas are the nodes' own
(`base-impl/arithmetic.lisp:57-103`, `matrix-ops.lisp:6-95`, `criterion.lisp:81-106`).

B200-first choices (SURVEY §7 hard parts e, f; §8e, §8f-1/2):
  * parameters and gradients live in ONE arena each, so Adam/SGD is one kernel over the bucket and
    the data-parallel gradient allreduce is one NCCL call;
  * every temporary comes from a StaticPool (the VMAllocation analogue), so forward + backward +
    allreduce + optimizer is a fixed launch sequence, captured once into a CUDA graph;
  * host scalars that change per step (Adam's lr_t) are read from device memory by the kernel.
"""
from __future__ import annotations

import os
import ctypes as C
import math
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import _lib, ops
from .tensor import CUDATensor, StaticPool, Storage, make_tensor, tensor_at, view

_P = _lib


# ------------------------------------------------------------------------------- tape
class Tape:
    """Reverse-mode record of one forward pass (`forward->reverse-mode`, vm/compile.lisp:110-158)."""

    def __init__(self):
        self.nodes: List[Callable[[], None]] = []
        self.touched: List[CUDATensor] = []

    def reset(self):
        for t in self.touched:
            t._has_grad = False
        self.nodes.clear()
        self.touched.clear()


_tape: Optional[Tape] = None

# Node-level fusion (what the backend's `on-finalizing-compiling` pass does to the instruction sequence,
# cl-waffe2_b200/lisp/fusion.lisp): LinearLayer's MatMul + bias add (+ the `!relu` after it) become one
# wf2cu_gemm_ex; the `!relu` backward rides on the dX GEMM that produces its incoming gradient;
# `(!sum (softmax-cross-entropy ...))` is one kernel forward and one backward.  FUSE = False issues one
# launch per node, exactly as the unfused instruction sequence would (both are tested against the oracle).
FUSE = True
# Independent backward instructions on side streams (wf2cu_branch_*): a layer's weight-gradient GEMM and
# bias-gradient reduction run beside the input-gradient GEMM that the next backward node waits for.
FUSE_SCE_GRAD = os.environ.get("WF2CU_FUSE_SCE_GRAD", "1") != "0"   # A/B switch, see softmax_cross_entropy_sum
_known_seed: Optional["CUDATensor"] = None      # set by CompiledStep._trace around the forward pass
CONCURRENT = True
CONSTANT_SEED = os.environ.get("WF2CU_CONSTANT_SEED", "1") != "0"   # A/B switch, see CompiledStep._trace
_open_branches: List[int] = []


def _join_branches() -> None:
    lib = _lib.load()
    for b in _open_branches:
        _lib.check(lib.wf2cu_branch_join(b))
    _open_branches.clear()


def _tracked(t: CUDATensor) -> bool:
    return bool(getattr(t, "requires_grad", False) or getattr(t, "_track", False))


def _mark(out: CUDATensor, *inputs: CUDATensor) -> bool:
    track = _tape is not None and any(_tracked(i) for i in inputs)
    out._track = track
    return track


def _accumulate(t: CUDATensor, compute: Callable[[CUDATensor, bool], None]) -> None:
    """Gradient adder (`vm/utils.lisp:109-142`): the first contribution is written (Move), later
    ones are added.  A tensor produced by a fused `... + !relu` node carries `_relu_out`: its incoming
    gradient is multiplied by (out > 0) — by the producing GEMM's epilogue when `compute` can do that
    (it then sets `_mask_done`), by one extra element-wise launch otherwise."""
    if not _tracked(t):
        return
    if t.grad is None:
        like = getattr(t, "_grad_like", None)       # e.g. Conv2D's channels-last result: same layout for its gradient
        t.grad = like(t) if like is not None else make_tensor(t.shape)
    acc = bool(getattr(t, "_has_grad", False))
    masked = getattr(t, "_relu_out", False)
    t._mask_done = False
    if masked and acc:
        tmp = make_tensor(t.shape)          # a second contribution: mask it on its own, then add
        compute(tmp, False)
        if not t._mask_done:
            _mask_mul(tmp, t)
        ops.add_(t.grad, tmp)
    else:
        compute(t.grad, acc)
        if masked and not t._mask_done:
            _mask_mul(t.grad, t)
    t._has_grad = True
    _tape.touched.append(t)


def _mask_mul(g: CUDATensor, out: CUDATensor) -> None:
    """g <- g * (out > 0 ? 1 : 0)"""
    prog = [(_P.P_LOAD, 0, 0, 0), (_P.P_IMM, 1, 0, ops._imm(0.0)), (_P.P_CMPSEL | (_P.CMP["gt"] << 8), 1, 0, 1),
            (_P.P_LOAD, 2, 1, 0), (_P.P_MUL, 1, 2, 1), (_P.P_STORE, 1, 1, 0)]
    ops.fused(prog, [out, g])


# ------------------------------------------------------------------------------- nodes
def linear(x: CUDATensor, weight: CUDATensor, bias: Optional[CUDATensor], fuse_relu: bool = False) -> CUDATensor:
    """`step-linear` (nn/regression.lisp:14-18): y = x @ W^T + b, W is (out in).
    fuse_relu: the `!relu` that follows is folded in (FUSE only); `out` is then the activation and its
    incoming gradient gets the (out > 0) mask (out > 0 <=> pre-activation > 0)."""
    if FUSE and x.dims() == 2:
        out = ops.matmul(x, ops.t(weight), bias=bias, relu=fuse_relu)
    else:
        assert not fuse_relu
        out = ops.matmul(x, ops.t(weight))
        if bias is not None:
            ops.add_(out, view(_as_row(bias), [":broadcast", out.shape[0]], True))
    if _mark(out, x, weight, *([bias] if bias is not None else [])):
        out._relu_out = bool(fuse_relu)

        def bw():
            dout = out.grad
            lib = _lib.load()

            # MatMulNode backward: dA = dout . B^T, dB = A^T . dout with B = (!t W)
            def param_grads():
                _accumulate(weight, lambda g, acc: _gemm_into(g, acc, ops.t(dout), x))
                if bias is not None:
                    # the (!flexible bias) broadcast view accumulates dout over the batch axis
                    def bias_grad(g, acc):
                        gb = view(_as_row(g), [":broadcast", dout.shape[0]], True)
                        _lib.check(lib.wf2cu_reduce_sum(C.byref(dout.lower()), C.byref(gb.lower()), 1 if acc else 0, 0.0))
                    _accumulate(bias, bias_grad)

            branch = 0
            if CONCURRENT and _tracked(x):
                branch = 1 + len(_open_branches) % 3
                if branch in _open_branches:          # more than 3 layers in flight: wait for the oldest user
                    _lib.check(lib.wf2cu_branch_join(branch))
                    _open_branches.remove(branch)
            if branch:
                # nobody but the optimizer waits for dW / db: side stream; dX stays on the main stream
                _lib.check(lib.wf2cu_branch_begin(branch))
                try:
                    param_grads()
                finally:
                    _lib.check(lib.wf2cu_branch_end(branch))
                _open_branches.append(branch)
                _accumulate(x, lambda g, acc: _gemm_into(g, acc, dout, weight, target=x))
            elif CONCURRENT and bias is not None:
                # first layer (no dX): the bias-gradient reduction runs beside the weight-gradient GEMM
                branch = 1 + len(_open_branches) % 3
                if branch in _open_branches:
                    _lib.check(lib.wf2cu_branch_join(branch))
                    _open_branches.remove(branch)

                def bias_grad1(g, acc):
                    gb = view(_as_row(g), [":broadcast", dout.shape[0]], True)
                    _lib.check(lib.wf2cu_reduce_sum(C.byref(dout.lower()), C.byref(gb.lower()), 1 if acc else 0, 0.0))
                _lib.check(lib.wf2cu_branch_begin(branch))
                try:
                    _accumulate(bias, bias_grad1)
                finally:
                    _lib.check(lib.wf2cu_branch_end(branch))
                _open_branches.append(branch)
                _accumulate(x, lambda g, acc: _gemm_into(g, acc, dout, weight, target=x))
                _accumulate(weight, lambda g, acc: _gemm_into(g, acc, ops.t(dout), x))
            else:
                _accumulate(x, lambda g, acc: _gemm_into(g, acc, dout, weight, target=x))
                param_grads()
        _tape.nodes.append(bw)
    return out


def _as_row(b: CUDATensor) -> CUDATensor:
    """(out) -> (1 out) without moving data (`!flexible` + rank-up, base-impl/fundamental.lisp)."""
    if b.dims() == 2:
        return b
    t = tensor_at(b.vec, b.lower().offset, [1, b.shape[0]])
    return t


def _gemm_into(g: CUDATensor, acc: bool, a, b, target: Optional[CUDATensor] = None) -> None:
    mask = target if (FUSE and target is not None and getattr(target, "_relu_out", False) and g.dims() == 2) else None
    if not acc:
        ops.matmul(a, b, out=g, relu_mask=mask)
    else:
        ops.add_(g, ops.matmul(a, b, relu_mask=mask))
    if mask is not None:
        target._mask_done = True


def relu(x: CUDATensor) -> CUDATensor:
    """`!relu` = x * (x > 0) (nn/activation.lisp:11-26); backward of the MulNode: dx = dout * mask."""
    out = ops.relu(x)
    if _mark(out, x):
        def bw():
            def f(g, acc):
                prog = [(_P.P_LOAD, 0, 0, 0), (_P.P_IMM, 1, 0, ops._imm(0.0)), (_P.P_CMPSEL | (_P.CMP["gt"] << 8), 1, 0, 1),
                        (_P.P_LOAD, 2, 1, 0), (_P.P_MUL, 1, 2, 1)]
                if acc:
                    prog += [(_P.P_LOAD, 3, 2, 0), (_P.P_ADD, 1, 3, 1)]
                prog += [(_P.P_STORE, 1, 2, 0)]
                ops.fused(prog, [x, out.grad, g])
            _accumulate(x, f)
        _tape.nodes.append(bw)
    return out


def sigmoid(x: CUDATensor) -> CUDATensor:
    """`!sigmoid` (nn/activation.lisp:43); backward folded to dout * s * (1 - s)."""
    out = ops.sigmoid(x)
    if _mark(out, x):
        def bw():
            def f(g, acc):
                prog = [(_P.P_LOAD, 0, 0, 0), (_P.P_IMM, 1, 0, ops._imm(1.0)), (_P.P_SUB, 1, 1, 0), (_P.P_MUL, 1, 0, 1),
                        (_P.P_LOAD, 2, 1, 0), (_P.P_MUL, 1, 2, 1)]
                if acc:
                    prog += [(_P.P_LOAD, 3, 2, 0), (_P.P_ADD, 1, 3, 1)]
                prog += [(_P.P_STORE, 1, 2, 0)]
                ops.fused(prog, [out, out.grad, g])
            _accumulate(x, f)
        _tape.nodes.append(bw)
    return out


def softmax_cross_entropy(x: CUDATensor, labels: CUDATensor, delta: float = 1e-7) -> CUDATensor:
    """`softmax-cross-entropy` (nn/criterion.lisp:164-188, Softmax-Cross-Entropy-Node :89-117):
    z = softmax(x, avoid-overflow nil); out = -1 * (labels * log(z + delta)).
    Backward (`sce-diff-static`, :81-87): dx = (dy * (z - labels)) / batch-size."""
    z = ops.softmax(x, avoid_overflow=False, axis=1)
    out = make_tensor(x.shape)
    prog = [(_P.P_LOAD, 0, 0, 0), (_P.P_IMM, 1, 0, ops._imm(delta)), (_P.P_ADD, 0, 0, 1),
            (_P.P_UNARY, 0, 0, _P.UNARY["loge"]), (_P.P_LOAD, 2, 1, 0), (_P.P_MUL, 0, 2, 0),
            (_P.P_IMM, 1, 0, ops._imm(-1.0)), (_P.P_MUL, 0, 1, 0), (_P.P_STORE, 0, 2, 0)]
    ops.fused(prog, [z, labels, out])
    if _mark(out, x):
        batch = float(x.shape[-2])

        def bw():
            def f(g, acc):
                p = [(_P.P_LOAD, 0, 0, 0), (_P.P_LOAD, 1, 1, 0), (_P.P_SUB, 0, 0, 1), (_P.P_LOAD, 2, 2, 0),
                     (_P.P_MUL, 0, 2, 0), (_P.P_IMM, 1, 0, ops._imm(batch)), (_P.P_DIV, 0, 0, 1)]
                if acc:
                    p += [(_P.P_LOAD, 3, 3, 0), (_P.P_ADD, 0, 3, 0)]
                p += [(_P.P_STORE, 0, 3, 0)]
                ops.fused(p, [z, labels, out.grad, g])
            _accumulate(x, f)
        _tape.nodes.append(bw)
    return out


def softmax_cross_entropy_sum(x: CUDATensor, labels: CUDATensor, delta: float = 1e-7) -> CUDATensor:
    """`(!sum (softmax-cross-entropy x labels))` as one kernel forward (z saved) and one backward
    (`sce-diff-static` with the summed loss' scalar gradient) — wf2cu_sce_sum_forward / wf2cu_sce_backward."""
    lib = _lib.load()
    rows, cols = x.shape[-2], x.shape[-1]
    if x.dims() != 2 or cols > 1024 or not (x.is_contiguous() and labels.is_contiguous()):
        return reduce_sum(softmax_cross_entropy(x, labels, delta))
    z = make_tensor(x.shape)
    out = make_tensor([1, 1])
    # A compiled training step seeds the summed loss with a constant it knows before the forward pass (_known_seed): the forward
    # launch then writes the logits' gradient as well — `sce-diff-static` with the same operations — and the backward node has
    # nothing to launch (checked there: first contribution, and the loss' incoming gradient IS that seed).
    seed = _known_seed if (FUSE_SCE_GRAD and _tape is not None and _tracked(x) and x.grad is None) else None
    if seed is not None:
        x.grad = make_tensor(x.shape)
        _lib.check(lib.wf2cu_sce_sum_forward_grad(x.vec.ptr, x.lower().offset, labels.vec.ptr, labels.lower().offset,
                                                  z.vec.ptr, z.lower().offset, out.vec.ptr, out.lower().offset, rows, cols, float(delta),
                                                  seed.vec.ptr, seed.lower().offset, x.grad.vec.ptr, x.grad.lower().offset))
    else:
        _lib.check(lib.wf2cu_sce_sum_forward(x.vec.ptr, x.lower().offset, labels.vec.ptr, labels.lower().offset,
                                             z.vec.ptr, z.lower().offset, out.vec.ptr, out.lower().offset, rows, cols, float(delta)))
    if _mark(out, x):
        def bw():
            dy = out.grad
            if seed is not None and dy is seed and not getattr(x, "_has_grad", False) and not getattr(x, "_relu_out", False):
                x._has_grad = True                   # written by the forward launch
                _tape.touched.append(x)
                return

            def f(g, acc):
                _lib.check(lib.wf2cu_sce_backward(z.vec.ptr, z.lower().offset, labels.vec.ptr, labels.lower().offset,
                                                  dy.vec.ptr, dy.lower().offset, g.vec.ptr, g.lower().offset, rows, cols, 1 if acc else 0))
            _accumulate(x, f)
        _tape.nodes.append(bw)
    return out


def reduce_sum(x: CUDATensor, mean: bool = False) -> CUDATensor:
    """`(!sum x)` / `(!mean x)` to a (1 .. 1) tensor; backward broadcasts dout (and divides by the
    count for mean: the `!div` node's dx = dout / n)."""
    out = ops.mean(x) if mean else ops.sum(x)
    if _mark(out, x):
        n = float(x.numel())

        def bw():
            gb = view(out.grad, *[[":broadcast", s] for s in x.shape])

            def f(g, acc):
                if mean:
                    tmp = ops.div(gb, n)
                    if acc:
                        ops.add_(g, tmp)
                    else:
                        ops.move_(g, tmp)
                elif acc:
                    ops.add_(g, gb)
                else:
                    ops.move_(g, gb)
            _accumulate(x, f)
        _tape.nodes.append(bw)
    return out


def mse(x: CUDATensor, y: CUDATensor, reduction: str = "mean") -> CUDATensor:
    """`mse` (nn/criterion.lisp:53-79): l = x - y; reduction(l * l).  Backward: MulNode gives
    (dout*l) to both operands (the same tensor l), SubNode gives (dl, -dl)."""
    l = ops.sub(x, y)
    sq = ops.mul(l, l)
    tracked = _mark(sq, x, y)
    if tracked:
        def bw():
            d = sq.grad

            def fx(g, acc):
                prog = [(_P.P_LOAD, 0, 0, 0), (_P.P_LOAD, 1, 1, 0), (_P.P_MUL, 0, 1, 0), (_P.P_ADD, 0, 0, 0)]
                if acc:
                    prog += [(_P.P_LOAD, 2, 2, 0), (_P.P_ADD, 0, 2, 0)]
                prog += [(_P.P_STORE, 0, 2, 0)]
                ops.fused(prog, [l, d, g])

            def fy(g, acc):
                prog = [(_P.P_LOAD, 0, 0, 0), (_P.P_LOAD, 1, 1, 0), (_P.P_MUL, 0, 1, 0), (_P.P_ADD, 0, 0, 0),
                        (_P.P_UNARY, 0, 0, _P.UNARY["neg"])]
                if acc:
                    prog += [(_P.P_LOAD, 2, 2, 0), (_P.P_ADD, 0, 2, 0)]
                prog += [(_P.P_STORE, 0, 2, 0)]
                ops.fused(prog, [l, d, g])
            _accumulate(x, fx)
            _accumulate(y, fy)
        _tape.nodes.append(bw)
    if reduction == "mean":
        return reduce_sum(sq, mean=True)
    if reduction == "sum":
        return reduce_sum(sq)
    return sq


# ------------------------------------------------------------------------------- parameters
def bucket_layout(shapes: Sequence[Sequence[int]]):
    """Element offsets of each parameter inside the flat bucket (every tensor 16-byte aligned) and the
    bucket length.  Pure host logic: the same layout is used for parameters, gradients and Adam state,
    and is what the NCCL allreduce sums over."""
    offsets = [0]
    for s in shapes:
        n = int(np.prod(s))
        offsets.append(offsets[-1] + ((n + 3) & ~3))
    return offsets[:-1], offsets[-1]


def shard_rows(global_batch: int, rank: int, world: int):
    """Rows [lo, hi) of the global batch owned by `rank` (batch sharding of SURVEY §8e; the reference
    expresses a minibatch the same way with `tensor-displace-to`, examples/mnist/mlp.lisp:46-48)."""
    if global_batch % world:
        raise ValueError(f"global batch {global_batch} is not divisible by world size {world}")
    per = global_batch // world
    return rank * per, (rank + 1) * per


class ParameterArena:
    """All parameters in one storage, all gradients in another (the allocator hook that
    makes a single gradient bucket possible)."""

    def __init__(self, shapes: Sequence[Sequence[int]]):
        offs, total = bucket_layout(shapes)
        self.offsets = offs + [total]
        self.total = total
        _lib.init(_lib._inited_device if _lib._inited_device is not None else 0)
        self.param_store = Storage(self.total * 4)
        self.grad_store = Storage(self.total * 4)
        self.params: List[CUDATensor] = []
        for shp, off in zip(shapes, self.offsets):
            p = tensor_at(self.param_store, off, shp, requires_grad=True)
            p.grad = tensor_at(self.grad_store, off, shp)
            self.params.append(p)


def xavier_uniform(rng: np.random.Generator, out_f: int, in_f: int) -> np.ndarray:
    """`xavier-uniform` (distributions/weights.lisp): U(-a, a), a = sqrt(6 / (in + out)).  Sampled on the
    host with a seeded numpy RNG (SBCL's RNG is not reproducible outside SBCL) and uploaded in bulk."""
    a = math.sqrt(6.0 / (in_f + out_f))
    return rng.uniform(-a, a, size=(out_f, in_f)).astype(np.float32)


def bias_uniform(rng: np.random.Generator, out_f: int) -> np.ndarray:
    """U(-k, k), k = sqrt(1 / out-features) (nn/regression.lisp:51-55)."""
    k = math.sqrt(1.0 / out_f)
    return rng.uniform(-k, k, size=(out_f,)).astype(np.float32)


def mlp_init_arrays(dims: Sequence[int], seed: int = 0):
    """Host arrays [(W, b), ...] for an MLP with layer widths `dims`."""
    rng = np.random.default_rng(seed)
    return [(xavier_uniform(rng, dims[i + 1], dims[i]), bias_uniform(rng, dims[i + 1])) for i in range(len(dims) - 1)]


class MLPSequence:
    """`MLP-Sequence` (examples/mnist/mlp.lisp:12-19) / the sin-wave MLP (examples/mlp_sin_wave.lisp:41-52):
    LinearLayer -> activation -> ... -> LinearLayer."""

    def __init__(self, dims: Sequence[int], activation: str = "relu", init=None, seed: int = 0):
        self.dims = list(dims)
        self.activation = {"relu": relu, "sigmoid": sigmoid}[activation]
        init = init if init is not None else mlp_init_arrays(dims, seed)
        shapes = []
        for i in range(len(dims) - 1):
            shapes += [(dims[i + 1], dims[i]), (dims[i + 1],)]
        self.arena = ParameterArena(shapes)
        for i, (w, b) in enumerate(init):
            self.arena.params[2 * i].copy_from_numpy(w)
            self.arena.params[2 * i + 1].copy_from_numpy(b)

    def parameters(self) -> List[CUDATensor]:
        return self.arena.params

    def named_parameters(self):
        """(composite, slot, tensor) in parameter order: LinearLayer's slots are `weights` and `bias`
        (nn/regression.lisp:20-55), which gives the state-dict keys `param:linearlayer.<nth>.weights`."""
        out = []
        for i in range(len(self.dims) - 1):
            out += [("LinearLayer", "weights", self.arena.params[2 * i]), ("LinearLayer", "bias", self.arena.params[2 * i + 1])]
        return out

    def __call__(self, x: CUDATensor) -> CUDATensor:
        p = self.arena.params
        n = len(self.dims) - 1
        for i in range(n):
            fuse = FUSE and i != n - 1 and self.activation is relu and x.dims() == 2
            x = linear(x, p[2 * i], p[2 * i + 1], fuse_relu=fuse)
            if i != n - 1 and not fuse:
                x = self.activation(x)
        return x


# ------------------------------------------------------------------------------- optimizers
class SGD:
    """`SGD` (optimizers/impls/sgd.lisp:4-35) over the whole bucket."""

    def __init__(self, arena: ParameterArena, lr: float = 1e-3):
        self.arena, self.lr = arena, float(lr)

    def pre_step(self):
        pass

    def step(self, grad_scale: float = 1.0):
        a = self.arena
        lib = _lib.load()
        if grad_scale != 1.0:
            g = tensor_at(a.grad_store, 0, [a.total])
            ops.mul_(g, grad_scale)
        _lib.check(lib.wf2cu_sgd_step(a.param_store.ptr, 0, a.grad_store.ptr, 0, a.total, self.lr))


class Adam:
    """`Adam` (optimizers/impls/adam.lisp:5-116) over the whole bucket; lr_t lives in device memory."""

    def __init__(self, arena: ParameterArena, lr=1e-3, eps=1e-7, beta1=0.9, beta2=0.999):
        self.arena = arena
        self.lr, self.eps, self.b1, self.b2 = (float(np.float32(v)) for v in (lr, eps, beta1, beta2))
        self.n = 0
        self.m = Storage(arena.total * 4)
        self.v = Storage(arena.total * 4)
        self.lr_t = Storage(4)

    def lr_at(self, n: int) -> float:
        """`adam-step-lr` (adam.lisp:81-94) in single precision."""
        f = np.float32
        b1n, b2n = f(self.b1) ** f(n), f(self.b2) ** f(n)
        return float(f(self.lr) * (np.sqrt(f(1.0) - f(b2n)) / (f(1.0) - f(b1n))))

    def pre_step(self):
        """Host part of a step: bump N and publish lr_t (4-byte H2D; outside the captured graph)."""
        self.n += 1
        _lib.check(_lib.load().wf2cu_set_f32(self.lr_t.ptr, 0, self.lr_at(self.n)))

    def step(self, grad_scale: float = 1.0):
        a = self.arena
        _lib.check(_lib.load().wf2cu_adam_step(a.param_store.ptr, 0, a.grad_store.ptr, 0, self.m.ptr, 0, self.v.ptr, 0,
                                               a.total, 0.0, self.lr_t.ptr, self.b1, self.b2, self.eps, grad_scale))

    # ---- state-dict slots (`find-params` on the optimizer, state-dicts.lisp:204-222): per parameter idx
    def _slice(self, store: Storage, idx: int) -> CUDATensor:
        a = self.arena
        return tensor_at(store, a.offsets[idx], a.params[idx].shape)

    def state_for(self, idx: int):
        """[(slot, lisp-type-name, value)] as `(type-of value)` prints them in SBCL, lower case."""
        int_type = "bit" if self.n <= 1 else "(integer 0 4611686018427387903)"
        return [("lr", "single-float", self.lr), ("eps", "single-float", self.eps), ("beta1", "single-float", self.b1),
                ("beta2", "single-float", self.b2), ("n", int_type, int(self.n)),
                ("m", "cudatensor", self._slice(self.m, idx)), ("v", "cudatensor", self._slice(self.v, idx))]

    def restore(self, idx: int, slot: str, value) -> None:
        """Scalars are shared by every parameter's optimizer instance in the reference; here they are one."""
        if slot == "n":
            self.n = int(value)
        elif slot in ("lr", "eps"):
            setattr(self, slot, float(np.float32(value)))
        elif slot in ("beta1", "beta2"):
            setattr(self, "b1" if slot == "beta1" else "b2", float(np.float32(value)))
        else:
            raise KeyError(f"Adam: no slot {slot!r}")

    def step_p2p(self):
        """Data-parallel step as ONE kernel: sum of all ranks' gradient arenas over NVLink peer memory,
        x 1/world, Adam on the local replica (csrc/p2p.cu).  Needs `p2p_attach` first."""
        a = self.arena
        _lib.check(_lib.load().wf2cu_p2p_allreduce_adam_step(a.param_store.ptr, 0, 0, self.m.ptr, 0, self.v.ptr, 0,
                                                             a.total, 0.0, self.lr_t.ptr, self.b1, self.b2, self.eps))


def p2p_attach(arena: ParameterArena, rank: int, world: int, all_gather_bytes) -> None:
    """Map every peer's gradient arena and signal page into this process (CUDA IPC).
    `all_gather_bytes(b: bytes) -> List[bytes]` is the host transport for the 64-byte handles
    (torch.distributed in bench.py; anything that can move 192 bytes per rank works)."""
    lib = _lib.load()
    hg = C.create_string_buffer(64)
    hp = C.create_string_buffer(64)
    hs = C.create_string_buffer(64)
    _lib.check(lib.wf2cu_p2p_export(arena.grad_store.ptr, hg))
    _lib.check(lib.wf2cu_p2p_export(arena.param_store.ptr, hp))
    _lib.check(lib.wf2cu_p2p_signal_export(hs))
    blobs = all_gather_bytes(hg.raw + hp.raw + hs.raw)
    if len(blobs) != world or any(len(b) != 192 for b in blobs):
        raise ValueError("p2p_attach: all_gather_bytes must return one 192-byte blob per rank")
    grads = b"".join(b[:64] for b in blobs)
    params = b"".join(b[64:128] for b in blobs)
    sigs = b"".join(b[128:] for b in blobs)
    _lib.check(lib.wf2cu_p2p_attach(rank, world, arena.grad_store.ptr, arena.param_store.ptr,
                                    C.create_string_buffer(grads, len(grads)), C.create_string_buffer(params, len(params)),
                                    C.create_string_buffer(sigs, len(sigs))))


def p2p_detach() -> None:
    _lib.check(_lib.load().wf2cu_p2p_detach())


def p2p_timed_out() -> bool:
    """The exchange kernel's sticky give-up flag: a mapped pinned host word, read without a device sync."""
    f = C.c_int(0)
    _lib.check(_lib.load().wf2cu_p2p_status(C.byref(f)))
    return bool(f.value)


# ------------------------------------------------------------------------------- compiled step
class CompiledStep:
    """`build` + `forward`/`backward`/`call-optimizer!` of the reference
    (vm/generic-tensor/acceptor.lisp:307-508) for one (model, loss, optimizer) triple with static input
    shapes: the launch sequence is traced once under a StaticPool and replayed as a CUDA graph.

    world > 1: batch-sharded data parallelism — every rank runs the same graph on its shard; the
    gradient bucket is summed with one NCCL allreduce inside the graph and scaled by 1/world in the
    optimizer kernel."""

    def __init__(self, model: MLPSequence, loss: str, optimizer, x_shape, y_shape, world: int = 1, use_graph: bool = True,
                 exchange: str = "nccl"):
        self.model, self.loss_kind, self.opt, self.world = model, loss, optimizer, int(world)
        if exchange not in ("nccl", "p2p"):
            raise ValueError(f"exchange must be 'nccl' or 'p2p', got {exchange!r}")
        self.exchange = exchange
        self.x = make_tensor(x_shape)
        self.y = make_tensor(y_shape)
        self.pool = StaticPool()
        self.tape = Tape()
        self.loss: Optional[CUDATensor] = None
        self.graph = None
        self.use_graph = use_graph
        self._runs = 0
        # measurement aid: run the peer-memory exchange kernel on a world of ONE attached rank (what the kernel itself costs
        # against the plain Adam launch, with nothing on the wire; scripts/p2p_world1.py)
        self.p2p_single = False
        self._seed = make_tensor([1, 1])
        ops.fill_(self._seed, 1.0)

    def _loss(self, out: CUDATensor) -> CUDATensor:
        if self.loss_kind == "softmax-cross-entropy-sum":      # examples/mnist/mlp.lisp:24-28
            if FUSE:
                return softmax_cross_entropy_sum(out, self.y)
            return reduce_sum(softmax_cross_entropy(out, self.y))
        if self.loss_kind == "mse-mean":                       # examples/mlp_sin_wave.lisp
            return mse(out, self.y, reduction="mean")
        raise ValueError(self.loss_kind)

    def _trace(self):
        """forward, backward, [allreduce], optimizer: all launches of one step."""
        global _tape
        lib = _lib.load()
        self.tape.reset()
        _tape = self.tape
        try:
            with self.pool:
                side_seed = None
                if CONCURRENT and not CONSTANT_SEED:
                    # WF2CU_CONSTANT_SEED=0: the seed refilled every step, on a side branch beside the forward pass
                    side_seed = make_tensor([1, 1])
                    _lib.check(lib.wf2cu_branch_begin(3))
                    try:
                        ops.fill_(side_seed, 1.0)
                    finally:
                        _lib.check(lib.wf2cu_branch_end(3))
                global _known_seed
                _known_seed = self._seed if (CONCURRENT and CONSTANT_SEED) else None
                try:
                    loss = self._loss(self.model(self.x))
                finally:
                    _known_seed = None
                if side_seed is not None:
                    _lib.check(lib.wf2cu_branch_join(3))
                if CONCURRENT and CONSTANT_SEED and tuple(loss.shape) == (1, 1):
                    # the backward seed (dL/dL = 1) is the same value every step and nothing writes it: it lives outside the step's
                    # pool, filled once.  (Refilled per step on a side branch it was one launch and two graph edges — a fork and a
                    # join on the main chain — of a step in which an edge costs about as much as a small kernel.)
                    loss.grad = self._seed
                elif side_seed is not None and tuple(loss.shape) == (1, 1):
                    loss.grad = side_seed
                else:
                    loss.grad = make_tensor(loss.shape)
                    ops.fill_(loss.grad, 1.0)
                loss._has_grad = True
                for node in reversed(self.tape.nodes):
                    node()
                _join_branches()               # every gradient is complete before the exchange / optimizer
                if self.exchange == "p2p" and (self.world > 1 or self.p2p_single):
                    self.opt.step_p2p()         # exchange + 1/world + Adam in one kernel over peer memory
                else:
                    if self.world > 1:
                        a = self.model.arena
                        _lib.check(lib.wf2cu_allreduce_sum_f32(a.grad_store.ptr, 0, a.total))
                    self.opt.step(1.0 / self.world if self.world > 1 else 1.0)
        finally:
            _tape = None
        self.loss = loss

    def set_inputs(self, x: np.ndarray, y: np.ndarray):
        self.x.copy_from_numpy(x)
        self.y.copy_from_numpy(y)

    def step(self):
        """One training step on the inputs currently in self.x / self.y (asynchronous)."""
        lib = _lib.load()
        self.opt.pre_step()
        if not self.use_graph or self._runs == 0:
            self._trace()                       # first run allocates the static plan
        else:
            if self.graph is None:
                self.pool.frozen = True
                ex = C.c_void_p()
                _lib.check(lib.wf2cu_graph_begin())
                try:
                    self._trace()
                finally:
                    rc = lib.wf2cu_graph_end(C.byref(ex))
                _lib.check(rc)
                self.graph = ex
            _lib.check(lib.wf2cu_graph_launch(self.graph))
        self._runs += 1
        if self.world > 1 and self.exchange == "p2p" and p2p_timed_out():
            # sticky and fatal (csrc/p2p.cu): a peer never arrived, every rank's exchange kernel now returns at entry
            raise RuntimeError("CompiledStep: the peer-memory gradient exchange timed out waiting for a rank; the replicas were "
                               "not updated from that step on — rebuild the process group and restore from a checkpoint")

    def close(self):
        """Destroy the captured graph (must happen before the communicator it references goes away)."""
        if self.graph is not None:
            _lib.check(_lib.load().wf2cu_sync())
            _lib.check(_lib.load().wf2cu_graph_destroy(self.graph))
            self.graph = None

    def loss_value(self) -> float:
        """`->scal` of the loss: one 4-byte D2H, the only sync of a step."""
        return self.loss.vref(self.loss.lower().offset)
