"""oracle/conv.py — TEST INFRASTRUCTURE, NOT PRODUCT.

CPU restatement of the reference's unfold / Conv2D / pooling path, node by node in the
order the reference composes it, on top of oracle/loopnest.py — so the leaves are the reference's own
C kernels (`waffe2_im2col_s_column`, `waffe2_col2im_s_column`, `waffe2_smax`, the arithmetic kernels)
and OpenBLAS sgemm:

  * padding                 `source/base-impl/utils.lisp:4-123`      move into a view of a larger tensor
  * Im2ColNode / Col2ImNode `source/base-impl/unfold.lisp:9-110`, CPU impl `source/backends/cpu/im2col.lisp:4-50`
  * !im2col / unfold        `source/nn/im2col.lisp:15-98`            (N C kh kw Ho Wo) -> permute -> (N*Ho*Wo, C*kh*kw)
  * Conv2D                  `source/nn/conv.lisp:88-105`             unfold @ (!t (reshape W)) + bias -> permute to NCHW
  * MaxPool2D / AvgPool2D   `source/nn/pool.lisp:73-105`             unfold -> (.., kh*kw) -> !max / !mean axis 1
  * MNIST-CNN               `examples/mnist/cnn.lisp:5-26`

Semantics kept as written: the image is zero-padded before Im2Col and the node then skips the last
`pad` output rows/columns (`im2col.c:41,43`: the loop bound subtracts pad); dilation only enters that
bound, never the sampled coordinate, and walks out of the buffer for dilation > 1 (`conv.lisp:41`:
"currently not working, please set 1") — dilation must be 1 here.

One point the reference leaves unpinned: in the backward pass `Col2ImNode` receives the gradient as a
*permuted view* (Permute-Node backward, `fundamental.lisp:937-956`) and the CPU impl hands the raw
storage pointer to C (`cpu/im2col.lisp:27-50`).  The reference's own tests only exercise H_out = W_out = 1
or element-uniform gradients (`nn/t/conv.lisp:43-52,84-110`), where reading the storage in either order
gives the same result.  This restatement, and the CUDA backend, index the gradient through its view
(the mathematically correct gradient).
"""
from __future__ import annotations

import math

import numpy as np

from . import loopnest as L

T = L.T


# ----------------------------------------------------------------------------- shapes
def conv_out_size(n, padding, dilation, kernel, stride):
    """`conv-out-size` (nn/pool.lisp:9-16)."""
    return int(math.floor(1 + (n + 2 * padding - dilation * (kernel - 1) - 1) / stride))


def pool_out_size(n, padding, kernel, stride):
    """`pool-out-size` (nn/pool.lisp:18-20)."""
    return int(math.floor(1 + (n + 2 * padding - kernel) / stride))


def _plain(t: L.OTensor) -> bool:
    return (not t.permuted_p()) and all(L.viewtype(v.view) == "t" for v in t.view) and t.offset == 0


def contiguous(t: L.OTensor) -> L.OTensor:
    """`->contiguous` (fundamental.lisp:1059-1107): a Move into a fresh tensor when permuted / viewed."""
    return t if _plain(t) else L.copy(t)


def reshape(t: L.OTensor, shape) -> L.OTensor:
    """`!reshape` (fundamental.lisp:392-425): `->contiguous`, then share the storage under the new shape."""
    t = contiguous(t)
    shape = list(shape)
    n = t.numel()
    if T in shape or -1 in shape:
        k = shape.index(T) if T in shape else shape.index(-1)
        rest = int(np.prod([s for i, s in enumerate(shape) if i != k]))
        shape[k] = n // rest
    assert int(np.prod(shape)) == n, (t.shape, shape)
    return L.OTensor(t.vec, shape, order=t.order)


def torch_order(t: L.OTensor, *orders):
    """`torch-order` (reshapers.lisp:41-54)."""
    d = t.dims() - 1
    return [d - o for o in orders]


# ----------------------------------------------------------------------------- padding
def padding(x: L.OTensor, pad_width):
    """`padding` (base-impl/utils.lisp:96-123).  The new tensor comes from `make-input` (pool memory, zero
    on first allocation): the border is taken as 0."""
    sizes = [s if w is T else s + w[0] + w[1] for s, w in zip(x.shape, pad_width)]
    if sizes == x.shape:
        return x
    new = L.make_tensor(sizes, order=x.order)
    win = L.view(new, *[T if w is T else [w[0], w[0] + s] for s, w in zip(x.shape, pad_width)])
    L.move_(win, x)
    return new


def padding_bw(dpadded: L.OTensor, x_shape, pad_width) -> L.OTensor:
    """ViewTensorNode + MoveTensorNode backward: the window of the padded gradient."""
    return L.view(dpadded, *[T if w is T else [w[0], w[0] + s] for s, w in zip(x_shape, pad_width)])


# ----------------------------------------------------------------------------- unfold
class UnfoldCtx:
    pass


def unfold_fw(x: L.OTensor, dilation, kernel_size, stride, pad):
    """`unfold` (nn/im2col.lisp:56-98): pad, Im2ColNode, permute (0 4 5 1 2 3), reshape to
    (N*Ho*Wo, C*kh*kw)."""
    assert tuple(dilation) == (1, 1), "dilation must be 1 (see module docstring)"
    N, C, Hin, Win = x.shape
    kh, kw = kernel_size
    Ho = conv_out_size(Hin, pad[0], dilation[0], kh, stride[0])
    Wo = conv_out_size(Win, pad[1], dilation[0], kw, stride[1])      # (car dilation) for both, as written (:90)
    ph, pw = Ho % stride[0], Wo % stride[1]
    pad_width = [T, T, (pad[0], pad[0] + ph), (pad[1], pad[1] + pw)]
    xp = contiguous(padding(x, pad_width))
    H, W = xp.shape[2], xp.shape[3]
    col = L.make_tensor([N, C, kh, kw, Ho, Wo], order=x.order)
    L._K().unfold("im2col", col.vec, N, C, H, W, Ho, Wo, kh, kw, pad[0], pad[1], stride[0], stride[1],
                  dilation[0], dilation[1], xp.vec)
    perm = L.permute(col, *torch_order(col, 0, 4, 5, 1, 2, 3))
    rows = reshape(perm, [N * Ho * Wo, C * kh * kw])
    ctx = UnfoldCtx()
    ctx.__dict__.update(N=N, C=C, H=H, W=W, Ho=Ho, Wo=Wo, kh=kh, kw=kw, pad=tuple(pad), stride=tuple(stride),
                        dilation=tuple(dilation), x_shape=list(x.shape), pad_width=pad_width)
    return rows, ctx


def unfold_bw(drows: L.OTensor, c: UnfoldCtx) -> L.OTensor:
    """Backward of the chain above: reshape back, permute back (a view), Col2ImNode, window of the
    padded gradient.  Returns a view with the input's shape."""
    d6 = reshape(drows, [c.N, c.Ho, c.Wo, c.C, c.kh, c.kw])
    dcol = L.permute(d6, *torch_order(d6, 0, 3, 4, 5, 1, 2))            # -> (N C kh kw Ho Wo), a view
    dcol = contiguous(dcol)                                             # indexed through its view (docstring)
    dimg = L.make_tensor([c.N, c.C, c.H, c.W], order=drows.order)
    L._K().unfold("col2im", dcol.vec, c.N, c.C, c.H, c.W, c.Ho, c.Wo, c.kh, c.kw, c.pad[0], c.pad[1],
                  c.stride[0], c.stride[1], c.dilation[0], c.dilation[1], dimg.vec)
    if [c.N, c.C, c.H, c.W] == c.x_shape:
        return dimg
    return padding_bw(dimg, c.x_shape, c.pad_width)


# ----------------------------------------------------------------------------- Conv2D
def _bcast_rows(b: L.OTensor, n: int) -> L.OTensor:
    row = L.OTensor(b.vec, [1, b.shape[-1]])
    return L.view(row, [":broadcast", n], T)


def conv2d_fw(x, W, b, stride=(1, 1), pad=(0, 0), dilation=(1, 1)):
    """`apply-conv2d` (nn/conv.lisp:88-105).  Returns the (N Cout Ho Wo) *permuted view* the reference
    returns, and the context for backward."""
    N = x.shape[0]
    Cout, _, kh, kw = W.shape
    rows, u = unfold_fw(x, dilation, (kh, kw), stride, pad)
    col_w = reshape(W, [Cout, T])
    out = L.matmul(rows, col_w, tb=True)
    if b is not None:
        L.arith_("add", out, _bcast_rows(b, out.shape[0]))
    o4 = reshape(out, [N, u.Ho, u.Wo, Cout])
    res = L.permute(o4, 3, 0, 2, 1)                                     # -> (N Cout Ho Wo)
    return res, (rows, col_w, u, W.shape, b is not None)


def conv2d_bw(dout: L.OTensor, ctx, need_dx=True):
    """dout has the output's shape (N Cout Ho Wo).  Permute-Node backward turns it into (N Ho Wo Cout);
    the reshape backward makes it contiguous; MatMulNode backward dA = dout.B^T, dB = A^T.dout;
    the bias's broadcast view accumulates dout over the rows."""
    rows, col_w, u, w_shape, has_bias = ctx
    Cout = w_shape[0]
    d4 = L.permute(dout, *torch_order(dout, 0, 2, 3, 1))               # (N Ho Wo Cout) view
    d2 = reshape(d4, [u.N * u.Ho * u.Wo, Cout])
    dWm = L.matmul(d2, rows, ta=True)                                   # (Cout, C*kh*kw)
    dW = reshape(dWm, list(w_shape))
    db = None
    if has_bias:
        db = L.make_tensor([1, Cout])
        L.arith_("add", L.view(db, [":broadcast", d2.shape[0]], T), d2)
    dx = None
    if need_dx:
        drows = L.matmul(d2, col_w)                                     # (M, C*kh*kw)
        dx = unfold_bw(drows, u)
    return dx, dW, db


# ----------------------------------------------------------------------------- pooling
def _pool_rows(x, kernel_size, stride, pad):
    N, C, Hin, Win = x.shape
    Ho = pool_out_size(Hin, pad[0], kernel_size[0], stride[0])
    Wo = pool_out_size(Win, pad[1], kernel_size[1], stride[1])
    rows, u = unfold_fw(x, (1, 1), kernel_size, stride, pad)
    win = reshape(rows, [T, kernel_size[0] * kernel_size[1]])
    return win, u, (N, C, Ho, Wo)


def maxpool2d_fw(x, kernel_size, stride=None, pad=(0, 0)):
    """`apply-maxpool2d` (nn/pool.lisp:73-88)."""
    stride = tuple(stride) if stride is not None else tuple(kernel_size)
    win, u, (N, C, Ho, Wo) = _pool_rows(x, kernel_size, stride, pad)
    m = L.maxmin("max", win)                                            # (M*C, 1)
    o4 = reshape(m, [N, Ho, Wo, C])
    return L.permute(o4, 3, 0, 2, 1), (win, m, u)


def maxpool2d_bw(dout, ctx):
    """MaxValue-Node backward (matrix-ops.lisp:311-334): mask = (a == max(a)) ; grad = mask * dout."""
    win, m, u = ctx
    d4 = L.permute(dout, *torch_order(dout, 0, 2, 3, 1))
    d = reshape(d4, [win.shape[0], 1])
    kk = win.shape[1]
    mask = L.compare("eq", win, L.view(m, T, [":broadcast", kk]))
    g = L.arith_("mul", mask, L.view(d, T, [":broadcast", kk]))
    drows = reshape(g, [u.N * u.Ho * u.Wo, u.C * u.kh * u.kw])
    return unfold_bw(drows, u)


def avgpool2d_fw(x, kernel_size, stride=None, pad=(0, 0)):
    """`apply-avgpool2d` (nn/pool.lisp:90-105)."""
    stride = tuple(stride) if stride is not None else tuple(kernel_size)
    win, u, (N, C, Ho, Wo) = _pool_rows(x, kernel_size, stride, pad)
    m = L.mean_(win, axis=1)
    o4 = reshape(m, [N, Ho, Wo, C])
    return L.permute(o4, 3, 0, 2, 1), (win, u)


def avgpool2d_bw(dout, ctx):
    """`!mean` backward: the `!div` node gives dout / count, the `!sum` broadcast spreads it."""
    win, u = ctx
    d4 = L.permute(dout, *torch_order(dout, 0, 2, 3, 1))
    d = reshape(d4, [win.shape[0], 1])
    kk = win.shape[1]
    d = L.scalar_op_("div", L.copy(d), float(np.float32(kk)))
    g = L.copy(L.view(d, T, [":broadcast", kk]))
    drows = reshape(g, [u.N * u.Ho * u.Wo, u.C * u.kh * u.kw])
    return unfold_bw(drows, u)


# ----------------------------------------------------------------------------- MNIST-CNN
class OracleCNN:
    """Conv2D(1,4,3x3) relu MaxPool(2) Conv2D(4,16,5x5) relu MaxPool(2) reshape LinearLayer(256,10);
    loss = (!sum (softmax-cross-entropy out y)); Adam (examples/mnist/cnn.lisp:5-26)."""

    def __init__(self, init, lr=1e-3, eps=1e-7, beta1=0.9, beta2=0.999):
        from . import mlp as M
        self.M = M
        self.P = []
        for w, b in init:
            self.P += [L.make_tensor(list(w.shape), data=w), L.make_tensor([len(b)], data=b)]
        f = np.float32
        self.lr, self.eps, self.b1, self.b2 = f(lr), f(eps), f(beta1), f(beta2)
        self.n = 0
        self.m = [np.zeros_like(p.vec) for p in self.P]
        self.v = [np.zeros_like(p.vec) for p in self.P]

    def forward_backward(self, X, Y):
        M = self.M
        W1, b1, W2, b2, W3, b3 = self.P
        x = L.make_tensor(list(X.shape), data=X)
        y = L.make_tensor(list(Y.shape), data=Y)
        c1, k1 = conv2d_fw(x, W1, b1)
        r1, m1 = M.relu_fw(c1)
        p1, q1 = maxpool2d_fw(r1, (2, 2))
        c2, k2 = conv2d_fw(p1, W2, b2)
        r2, m2 = M.relu_fw(c2)
        p2, q2 = maxpool2d_fw(r2, (2, 2))
        flat = reshape(p2, [X.shape[0], W3.shape[1]])
        out = M.linear_fw(flat, W3, b3)
        lmat, z = M.sce_fw(out, y)
        loss = L.sum_(lmat).vec[0]
        dy = L.make_tensor(out.shape, data=np.ones(out.shape, np.float32))
        d = M.sce_bw(dy, z, y)
        dflat, gW3, gb3 = M.linear_bw(d, flat, W3)
        dp2 = reshape(dflat, p2.shape)
        dr2 = maxpool2d_bw(dp2, q2)
        dc2 = L.arith_("mul", L.copy(dr2), m2)
        dp1, gW2, gb2 = conv2d_bw(dc2, k2)
        dr1 = maxpool2d_bw(dp1, q1)
        dc1 = L.arith_("mul", L.copy(dr1), m1)
        _, gW1, gb1 = conv2d_bw(dc1, k1, need_dx=False)
        self.out = out
        return float(loss), [gW1, gb1, gW2, gb2, gW3, gb3]

    def apply(self, grads):
        f = np.float32
        self.n += 1
        n = self.n
        lr_t = f(self.lr * (np.sqrt(f(1.0) - self.b2 ** f(n)) / (f(1.0) - self.b1 ** f(n))))
        for k, (p, g) in enumerate(zip(self.P, grads)):
            g = g.vec
            m, v = self.m[k], self.v[k]
            m[:] = m + (g - m) * (f(1.0) - self.b1)
            v[:] = v + (f(1.0) - self.b2) * (g * g - v)
            p.vec[:] = p.vec - (m * lr_t) / (self.eps + np.sqrt(v))

    def step(self, X, Y):
        loss, grads = self.forward_backward(X, Y)
        self.apply(grads)
        return loss
