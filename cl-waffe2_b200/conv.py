"""Unfold / Conv2D / pooling on CUDATensor — the host side of `csrc/conv.cu`.

Mirrors, with the reference's names and argument meaning:
  * `Im2ColNode` / `Col2ImNode` impls   `/root/reference/source/backends/cpu/im2col.lisp:4-50`      -> `im2col`, `col2im`
  * `padding`                            `source/base-impl/utils.lisp:4-123`                        -> `padding`
  * `unfold` / `!im2col`                 `source/nn/im2col.lisp:15-98`                              -> `unfold`
  * `Conv2D` (`apply-conv2d`)            `source/nn/conv.lisp:12-105`                               -> `conv2d`
  * `MaxPool2D` / `AvgPool2D`            `source/nn/pool.lisp:22-105`                               -> `maxpool2d`, `avgpool2d`
  * `MNIST-CNN`                          `examples/mnist/cnn.lisp:5-26`                             -> `MNISTCNN`

Two instruction sequences produce the same tensors (both are tested against the oracle):
  nn.FUSE = False  one launch per node, exactly the chain the reference composes (Move into the padded
                   view, Im2ColNode, permute + `->contiguous`, MatMul, bias add, ... Col2ImNode);
  nn.FUSE = True   what the backend's `on-finalizing-compiling` pass emits (lisp/fusion.lisp): `wf2cu_unfold`
                   / `wf2cu_fold` for the pad+im2col+permute+copy chain and its backward, the bias add and
                   `!relu` in the GEMM epilogue, `wf2cu_pool2d(_backward)` for a whole pooling layer.
Dilation must be 1 (the reference's unfold walks out of its buffer otherwise, nn/conv.lisp:41).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib, nn, ops
from .tensor import CUDATensor, T, as_strided, make_tensor, permute, reshape_contiguous, tensor_at, view

POOL_MAX, POOL_AVG = 0, 1


def _pair(v, name) -> Tuple[int, int]:
    """`maybe-tuple` (nn/conv.lisp:3-10)."""
    if isinstance(v, (int, np.integer)):
        return int(v), int(v)
    if isinstance(v, (list, tuple)) and len(v) == 2:
        return int(v[0]), int(v[1])
    raise ValueError(f"Conv2D: {name} should be given as fixnum or list(with length=2), but got {v!r}")


def conv_out_size(n, padding, dilation, kernel, stride) -> int:
    """`conv-out-size` (nn/pool.lisp:9-16)."""
    return int(math.floor(1 + (n + 2 * padding - dilation * (kernel - 1) - 1) / stride))


def pool_out_size(n, padding, kernel, stride) -> int:
    """`pool-out-size` (nn/pool.lisp:18-20)."""
    return int(math.floor(1 + (n + 2 * padding - kernel) / stride))


def torch_order(t: CUDATensor, *orders) -> List[int]:
    """`torch-order` (base-impl/reshapers.lisp:41-54)."""
    d = t.dims() - 1
    return [d - o for o in orders]


def reshape(t: CUDATensor, shape) -> CUDATensor:
    """`!reshape` (fundamental.lisp:392-425): `->contiguous` first, then the same storage under a new shape."""
    return reshape_contiguous(ops.contiguous(t), [True if s is T else s for s in shape])


class Geometry:
    """Slots of Im2ColNode (base-impl/unfold.lisp:9-23) for one unfold of an (N C Hin Win) input."""

    def __init__(self, x_shape, kernel_size, stride, pad, dilation=(1, 1)):
        self.N, self.C, self.Hin, self.Win = [int(s) for s in x_shape]
        self.kh, self.kw = kernel_size
        self.stride, self.pad, self.dilation = tuple(stride), tuple(pad), tuple(dilation)
        self.Ho = conv_out_size(self.Hin, pad[0], dilation[0], self.kh, stride[0])
        self.Wo = conv_out_size(self.Win, pad[1], dilation[0], self.kw, stride[1])   # (car dilation) twice, im2col.lisp:90
        # padding ((pad, pad + Ho mod stride) ...), im2col.lisp:91-94
        self.pad_width = [T, T, (pad[0], pad[0] + self.Ho % stride[0]), (pad[1], pad[1] + self.Wo % stride[1])]
        self.H = self.Hin + 2 * pad[0] + self.Ho % stride[0]
        self.W = self.Win + 2 * pad[1] + self.Wo % stride[1]

    def struct(self, padded: bool) -> _lib.UnfoldGeom:
        g = _lib.UnfoldGeom()
        g.N, g.C = self.N, self.C
        g.H, g.W = (self.H, self.W) if padded else (self.Hin, self.Win)
        g.Ho, g.Wo, g.KH, g.KW = self.Ho, self.Wo, self.kh, self.kw
        g.pad_h, g.pad_w = self.pad
        g.stride_h, g.stride_w = self.stride
        g.dilation_h, g.dilation_w = self.dilation
        return g

    @property
    def M(self):
        return self.N * self.Ho * self.Wo

    @property
    def K(self):
        return self.C * self.kh * self.kw


# ------------------------------------------------------------------------------- node-level ops
def im2col(x_padded: CUDATensor, col: CUDATensor, geom: Geometry) -> CUDATensor:
    """Im2ColNode :forward (backends/cpu/im2col.lisp:4-25): col (N C kh kw Ho Wo) <- padded image."""
    if not (x_padded.is_contiguous() and col.is_contiguous()):
        raise ValueError("Im2ColNode: both tensors must be dense (the reference hands raw pointers to C)")
    g = geom.struct(True)
    _lib.check(_lib.load().wf2cu_im2col(col.vec.ptr, col.lower().offset, x_padded.vec.ptr, x_padded.lower().offset, C.byref(g)))
    return col


def col2im(col: CUDATensor, img: CUDATensor, geom: Geometry) -> CUDATensor:
    """Col2ImNode :forward (backends/cpu/im2col.lisp:27-50): img <- 0 ; img += col."""
    if not (img.is_contiguous() and col.is_contiguous()):
        raise ValueError("Col2ImNode: both tensors must be dense")
    g = geom.struct(True)
    _lib.check(_lib.load().wf2cu_col2im(col.vec.ptr, col.lower().offset, img.vec.ptr, img.lower().offset, C.byref(g)))
    return img


def padding(x: CUDATensor, pad_width) -> CUDATensor:
    """`padding` (base-impl/utils.lisp:96-123): a zero tensor of the padded shape, `x` moved into its
    window.  (Taken from the pool, so the border is cleared explicitly each time.)"""
    sizes = [s if w is T else s + w[0] + w[1] for s, w in zip(x.shape, pad_width)]
    if sizes == x.shape:
        return x
    new = make_tensor(sizes)
    ops.fill_(new, 0.0)
    win = view(new, *[T if w is T else [w[0], w[0] + s] for s, w in zip(x.shape, pad_width)])
    ops.move_(win, x)
    return new


def _unfold_nodes(x: CUDATensor, geom: Geometry) -> CUDATensor:
    """The reference's chain, one launch per node."""
    xp = ops.contiguous(padding(x, geom.pad_width))
    col = make_tensor([geom.N, geom.C, geom.kh, geom.kw, geom.Ho, geom.Wo])
    if geom.pad != (0, 0):
        ops.fill_(col, 0.0)           # rows/columns the node never visits: pool memory is not re-zeroed
    im2col(xp, col, geom)
    perm = permute(col, *torch_order(col, 0, 4, 5, 1, 2, 3))
    return reshape(perm, [geom.M, geom.K])


def _fold_nodes(drows: CUDATensor, geom: Geometry) -> CUDATensor:
    """Backward of the chain, one launch per node; returns a view with the input's shape."""
    d6 = reshape(drows, [geom.N, geom.Ho, geom.Wo, geom.C, geom.kh, geom.kw])
    dcol = ops.contiguous(permute(d6, *torch_order(d6, 0, 3, 4, 5, 1, 2)))
    dimg = make_tensor([geom.N, geom.C, geom.H, geom.W])
    col2im(dcol, dimg, geom)
    if (geom.H, geom.W) == (geom.Hin, geom.Win):
        return dimg
    return view(dimg, *[T if w is T else [w[0], w[0] + s]
                        for s, w in zip([geom.N, geom.C, geom.Hin, geom.Win], geom.pad_width)])


def unfold(x: CUDATensor, dilation, kernel_size, stride, padding_) -> Tuple[CUDATensor, Geometry]:
    """`unfold` (nn/im2col.lisp:56-98): (N C H W) -> (N*Ho*Wo, C*kh*kw)."""
    if x.dims() != 4:
        raise ValueError("unfold: input must be a 4D tensor (N C H W)")
    geom = Geometry(x.shape, _pair(kernel_size, "kernel-size"), _pair(stride, "stride"), _pair(padding_, "padding"),
                    _pair(dilation, "dilation"))
    if nn.FUSE:
        rows = make_tensor([geom.M, geom.K])
        g = geom.struct(False)
        _lib.check(_lib.load().wf2cu_unfold(C.byref(x.lower()), rows.vec.ptr, rows.lower().offset, C.byref(g)))
    else:
        if geom.dilation != (1, 1):
            raise _lib.Wf2CudaError(3, "unfold: dilation must be 1")
        rows = _unfold_nodes(x, geom)
    return rows, geom


def _fold_into(drows: CUDATensor, geom: Geometry, g: CUDATensor, acc: bool) -> None:
    """Gradient of `unfold`'s input, written (or added) into g."""
    if nn.FUSE:
        gs = geom.struct(False)
        _lib.check(_lib.load().wf2cu_fold(drows.vec.ptr, drows.lower().offset, C.byref(g.lower()), C.byref(gs), 1 if acc else 0))
    else:
        d = _fold_nodes(drows, geom)
        if acc:
            ops.add_(g, d)
        else:
            ops.move_(g, d)


# ------------------------------------------------------------------------------- layout helpers
def _channels_last(n, c, h, w) -> CUDATensor:
    """A fresh (n c h w) tensor whose storage is (n h w c) — the layout `(!permute (!reshape out N Ho Wo C) 3 0 2 1)`
    leaves (nn/conv.lisp:105, nn/pool.lisp:88)."""
    base = make_tensor([n, h, w, c])
    return permute(base, 3, 0, 2, 1)


def _grad_like(t: CUDATensor) -> CUDATensor:
    if getattr(t, "_channels_last", False):
        return _channels_last(*t.shape)
    return make_tensor(t.shape)


def _rows_of(t: CUDATensor, rows: int, cols: int) -> Optional[CUDATensor]:
    """The (N*H*W, C) matrix under a channels-last (N C H W) tensor, without moving data; None if `t`
    is laid out differently."""
    n, c, h, w = t.shape
    if t.stride == [h * w * c, 1, w * c, c]:
        return tensor_at(t.vec, t.lower().offset, [rows, cols])
    return None


# ------------------------------------------------------------------------------- Conv2D
def conv2d(x: CUDATensor, weight: CUDATensor, bias: Optional[CUDATensor], stride=1, padding_=0, dilation=1,
           fuse_relu: bool = False) -> CUDATensor:
    """`apply-conv2d` (nn/conv.lisp:88-105): col = unfold(x); out = col @ (!t (reshape W Cout t)) + bias;
    (!permute (!reshape out N Ho Wo Cout) 3 0 2 1).  Returns the (N Cout Ho Wo) permuted view over
    (N Ho Wo Cout) storage, as the reference does.  fuse_relu (FUSE only): the `!relu` that follows is folded
    into the GEMM epilogue; the result then carries the relu-backward mask like nn.linear."""
    lib = _lib.load()
    cout, _, kh, kw = weight.shape
    rows, geom = unfold(x, dilation, (kh, kw), stride, padding_)
    wmat = reshape(weight, [cout, T])
    if nn.FUSE:
        out = ops.matmul(rows, ops.t(wmat), bias=bias, relu=fuse_relu)
    else:
        assert not fuse_relu
        out = ops.matmul(rows, ops.t(wmat))
        if bias is not None:
            ops.add_(out, view(nn._as_row(bias), [":broadcast", out.shape[0]], True))
    res = permute(reshape_contiguous(out, [geom.N, geom.Ho, geom.Wo, cout]), 3, 0, 2, 1)
    res._channels_last = True
    if nn._mark(res, x, weight, *([bias] if bias is not None else [])):
        res._relu_out = bool(fuse_relu)
        res._grad_like = _grad_like

        def bw():
            # Permute-Node backward + ReshapeTensorNode backward: (N Ho Wo Cout) rows of the incoming gradient
            d2 = _rows_of(res.grad, geom.M, cout)
            if d2 is None:
                d4 = permute(res.grad, *torch_order(res.grad, 0, 2, 3, 1))
                d2 = reshape(d4, [geom.M, cout])

            def param_grads():
                nn._accumulate(weight, lambda g, acc: nn._gemm_into(reshape_contiguous(g, [cout, geom.K]), acc, ops.t(d2), rows))
                if bias is not None:
                    def bias_grad(g, acc):
                        gb = view(nn._as_row(g), [":broadcast", d2.shape[0]], True)
                        _lib.check(lib.wf2cu_reduce_sum(C.byref(d2.lower()), C.byref(gb.lower()), 1 if acc else 0, 0.0))
                    nn._accumulate(bias, bias_grad)

            branch = 0
            if nn.CONCURRENT and nn._tracked(x):
                branch = 1 + len(nn._open_branches) % 3
                if branch in nn._open_branches:
                    _lib.check(lib.wf2cu_branch_join(branch))
                    nn._open_branches.remove(branch)
            if branch:
                _lib.check(lib.wf2cu_branch_begin(branch))
                try:
                    param_grads()
                finally:
                    _lib.check(lib.wf2cu_branch_end(branch))
                nn._open_branches.append(branch)
            else:
                param_grads()
            if nn._tracked(x):
                drows = ops.matmul(d2, wmat)                    # dA = dout . B^T with B = (!t wmat)
                nn._accumulate(x, lambda g, acc: _fold_into(drows, geom, g, acc))
        nn._tape.nodes.append(bw)
    return res


# ------------------------------------------------------------------------------- pooling
def _pool2d(mode: int, x: CUDATensor, kernel_size, stride, padding_) -> CUDATensor:
    if x.dims() != 4:
        raise ValueError("Pool2D: input must be a 4D tensor (N C H W)")
    ks = _pair(kernel_size, "kernel-size")
    st = _pair(stride if stride is not None else ks, "stride")
    pd = _pair(padding_, "padding")
    geom = Geometry(x.shape, ks, st, pd)
    geom.Ho = pool_out_size(geom.Hin, pd[0], ks[0], st[0])       # nn/pool.lisp:77-78 (equal to conv-out-size at dilation 1)
    geom.Wo = pool_out_size(geom.Win, pd[1], ks[1], st[1])
    lib = _lib.load()
    n, c = geom.N, geom.C
    if nn.FUSE:
        res = _channels_last(n, c, geom.Ho, geom.Wo)
        g = geom.struct(False)
        _lib.check(lib.wf2cu_pool2d(mode, C.byref(x.lower()), C.byref(res.lower()), C.byref(g)))
        win = m = None
    else:
        rows = _unfold_nodes(x, geom)
        win = reshape(rows, [T, ks[0] * ks[1]])
        m = ops.max(win) if mode == POOL_MAX else ops.mean(win, axis=1)
        res = permute(reshape(m, [n, geom.Ho, geom.Wo, c]), 3, 0, 2, 1)
    res._channels_last = True
    if nn._mark(res, x):
        res._grad_like = _grad_like

        def bw():
            def f(gx, acc):
                if nn.FUSE:
                    gs = geom.struct(False)
                    masked = bool(getattr(x, "_relu_out", False))      # x = relu(...) folded into its producer: mask here
                    _lib.check(lib.wf2cu_pool2d_backward(mode, C.byref(x.lower()), C.byref(res.lower()), C.byref(res.grad.lower()),
                                                         C.byref(gx.lower()), C.byref(gs), 1 if acc else 0, 1 if masked else 0))
                    if masked:
                        x._mask_done = True
                    return
                kk = ks[0] * ks[1]
                d4 = permute(res.grad, *torch_order(res.grad, 0, 2, 3, 1))
                d = reshape(d4, [win.shape[0], 1])
                db = view(d, T, [":broadcast", kk])
                if mode == POOL_MAX:
                    # MaxValue-Node backward (matrix-ops.lisp:311-334): (A=B a (max a)) * dout
                    mask = ops.compare(win, "eq", view(m, T, [":broadcast", kk]))
                    gwin = ops.mul_(mask, db)
                else:
                    gwin = ops.copy(view(ops.div(d, float(kk)), T, [":broadcast", kk]))
                drows = reshape(gwin, [geom.M, geom.K])
                dd = _fold_nodes(drows, geom)
                if acc:
                    ops.add_(gx, dd)
                else:
                    ops.move_(gx, dd)
            nn._accumulate(x, f)
        nn._tape.nodes.append(bw)
    return res


def maxpool2d(x: CUDATensor, kernel_size, stride=None, padding_=0) -> CUDATensor:
    """`MaxPool2D` (nn/pool.lisp:22-46, 73-88)."""
    return _pool2d(POOL_MAX, x, kernel_size, stride, padding_)


def avgpool2d(x: CUDATensor, kernel_size, stride=None, padding_=0) -> CUDATensor:
    """`AvgPool2D` (nn/pool.lisp:48-71, 90-105)."""
    return _pool2d(POOL_AVG, x, kernel_size, stride, padding_)


def flatten(x: CUDATensor, shape) -> CUDATensor:
    """`(asnode #'!reshape t n)` (examples/mnist/cnn.lisp:14): ReshapeTensorNode after `->contiguous`;
    backward reshapes dout back (fundamental.lisp:273-281)."""
    out = reshape(x, shape)
    if nn._mark(out, x):
        def bw():
            d = reshape_contiguous(out.grad, x.shape)

            def f(g, acc):
                if acc:
                    ops.add_(g, d)
                else:
                    ops.move_(g, d)
            nn._accumulate(x, f)
        nn._tape.nodes.append(bw)
    return out


# ------------------------------------------------------------------------------- MNIST-CNN
def conv_init_arrays(rng: np.random.Generator, in_c: int, out_c: int, ksize, groups: int = 1):
    """Conv2D's initialiser (nn/conv.lisp:72-86): W (out in/groups kh kw) and b (out) ~ U(-sqrt(k), sqrt(k)),
    k = groups / (in * kh * kw); sampled on the host with a seeded RNG and uploaded in bulk."""
    k = groups / (in_c * ksize[0] * ksize[1])
    a = math.sqrt(k)
    W = rng.uniform(-a, a, size=(out_c, in_c // groups, ksize[0], ksize[1])).astype(np.float32)
    b = rng.uniform(-a, a, size=(out_c,)).astype(np.float32)
    return W, b


def cnn_init_arrays(seed: int = 0, c1: int = 4, c2: int = 16):
    rng = np.random.default_rng(seed)
    W1, b1 = conv_init_arrays(rng, 1, c1, (3, 3))
    W2, b2 = conv_init_arrays(rng, c1, c2, (5, 5))
    W3 = nn.xavier_uniform(rng, 10, c2 * 4 * 4)
    b3 = nn.bias_uniform(rng, 10)
    return [(W1, b1), (W2, b2), (W3, b3)]


class MNISTCNN:
    """`MNIST-CNN` (examples/mnist/cnn.lisp:5-16): Conv2D(1, 4, 3x3) relu MaxPool2D(2) Conv2D(4, 16, 5x5) relu
    MaxPool2D(2) reshape LinearLayer(256, 10), on (N 1 28 28) inputs."""

    def __init__(self, out_channels1: int = 4, out_channels2: int = 16, init=None, seed: int = 0):
        self.c1, self.c2 = out_channels1, out_channels2
        init = init if init is not None else cnn_init_arrays(seed, out_channels1, out_channels2)
        shapes = []
        for w, b in init:
            shapes += [tuple(w.shape), tuple(b.shape)]
        self.arena = nn.ParameterArena(shapes)
        for i, (w, b) in enumerate(init):
            self.arena.params[2 * i].copy_from_numpy(w)
            self.arena.params[2 * i + 1].copy_from_numpy(b)

    def parameters(self) -> List[CUDATensor]:
        return self.arena.params

    def named_parameters(self):
        """Conv2D's slots are `weight` and `bias` (nn/conv.lisp:52-58), LinearLayer's `weights` and `bias`."""
        p = self.arena.params
        return [("Conv2D", "weight", p[0]), ("Conv2D", "bias", p[1]), ("Conv2D", "weight", p[2]), ("Conv2D", "bias", p[3]),
                ("LinearLayer", "weights", p[4]), ("LinearLayer", "bias", p[5])]

    def __call__(self, x: CUDATensor) -> CUDATensor:
        W1, b1, W2, b2, W3, b3 = self.arena.params
        h = conv2d(x, W1, b1, fuse_relu=nn.FUSE)
        if not nn.FUSE:
            h = nn.relu(h)
        h = maxpool2d(h, (2, 2))
        h = conv2d(h, W2, b2, fuse_relu=nn.FUSE)
        if not nn.FUSE:
            h = nn.relu(h)
        h = maxpool2d(h, (2, 2))
        h = flatten(h, [x.shape[0], self.c2 * 4 * 4])
        return nn.linear(h, W3, b3)
