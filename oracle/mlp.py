"""oracle/mlp.py — TEST INFRASTRUCTURE, NOT PRODUCT.

CPU restatement of the MLP training step of the reference's examples, node by node in the order the
reference composes them, on top of oracle/loopnest.py (so every leaf is the reference's own C
kernel / OpenBLAS sgemm):

  * LinearLayer           `source/nn/regression.lisp:14-18`   y = x @ (!t W) + b
  * !relu / !sigmoid      `source/nn/activation.lisp:11-43`
  * softmax-cross-entropy `source/nn/criterion.lisp:81-188`   (avoid-overflow nil, delta 1e-7)
  * mse                   `source/nn/criterion.lisp:53-79`
  * SGD / Adam            `source/optimizers/impls/{sgd,adam}.lisp`
  * examples              `examples/mnist/mlp.lisp:12-36`, `examples/mlp_sin_wave.lisp:41-78`

Backward passes follow the nodes' :backward definitions (`base-impl/arithmetic.lisp:57-103`,
`matrix-ops.lisp:6-95`, `criterion.lisp:81-106`); gradient accumulation order is not modelled
beyond that (a parameter receives exactly one contribution per step in these models).
"""
from __future__ import annotations

import numpy as np

from . import loopnest as L

T = L.T


def _bcast_row(b: L.OTensor, n: int) -> L.OTensor:
    """(!flexible bias) against a (n out) matrix: rank-up + (:broadcast n) on the new axis."""
    row = L.OTensor(b.vec, [1, b.shape[-1]])
    return L.view(row, [":broadcast", n], T)


def linear_fw(x, W, b):
    out = L.matmul(x, W, tb=True)
    return L.arith_("add", out, _bcast_row(b, out.shape[0]))


def linear_bw(dout, x, W):
    dx = L.matmul(dout, W)                    # dout . (W^T)^T
    dW = L.matmul(dout, x, ta=True)           # (x^T . dout)^T accumulated into the (out in) parameter
    db = L.make_tensor([1, dout.shape[1]])
    L.arith_("add", L.view(db, [":broadcast", dout.shape[0]], T), dout)   # sequential fp32 over the batch
    return dx, dW, db


def relu_fw(x):
    mask = L.where_scal("gt", x, 0.0)
    return L.arith_("mul", L.copy(x), mask), mask


def sigmoid_fw(x):
    e = L.scalar_op_("mul", L.copy(x), -1.0)
    e = L.unary("exp", e)
    e = L.scalar_op_("add", e, 1.0)
    one = L.make_tensor(x.shape, data=np.ones(x.shape, np.float32))
    return L.arith_("div", one, e)


def sigmoid_bw(dout, s):
    # chain of DivNode/AddNode/ExpNode/MulNode backward == dout * s * (1 - s)
    one = L.make_tensor(s.shape, data=np.ones(s.shape, np.float32))
    t = L.arith_("sub", one, s)
    t = L.arith_("mul", t, s)
    return L.arith_("mul", t, dout)


def sce_fw(x, labels, delta=1e-7):
    z = L.softmax(x, avoid_overflow=False, axis=1)
    t = L.scalar_op_("add", L.copy(z), delta)
    t = L.unary("log", t)
    t = L.arith_("mul", L.copy(labels), t)
    return L.scalar_op_("mul", t, -1.0), z


def sce_bw(dy, z, labels):
    t = L.arith_("sub", L.copy(z), labels)
    t = L.arith_("mul", L.copy(dy), t)
    return L.scalar_op_("div", t, float(z.shape[-2]))


class OracleMLP:
    def __init__(self, dims, activation, init, loss, opt="adam", lr=1e-3, eps=1e-7, beta1=0.9, beta2=0.999):
        self.dims, self.act, self.loss_kind, self.opt = list(dims), activation, loss, opt
        self.W = [L.make_tensor(list(w.shape), data=w) for w, _ in init]
        self.b = [L.make_tensor([len(b)], data=b) for _, b in init]
        f = np.float32
        self.lr, self.eps, self.b1, self.b2 = f(lr), f(eps), f(beta1), f(beta2)
        self.n = 0
        self.m = [np.zeros_like(p.vec) for p in self.W + self.b]
        self.v = [np.zeros_like(p.vec) for p in self.W + self.b]

    def params(self):
        return self.W + self.b

    def forward_backward(self, X: np.ndarray, Y: np.ndarray):
        x = L.make_tensor(list(X.shape), data=X)
        y = L.make_tensor(list(Y.shape), data=Y)
        acts, pre = [x], []
        h = x
        nl = len(self.W)
        for i in range(nl):
            h = linear_fw(h, self.W[i], self.b[i])
            if i != nl - 1:
                if self.act == "relu":
                    pre.append(h)
                    h, mask = relu_fw(h)
                    acts.append((h, mask))
                else:
                    h = sigmoid_fw(h)
                    acts.append((h, None))
        out = h
        if self.loss_kind == "softmax-cross-entropy-sum":
            lmat, z = sce_fw(out, y)
            loss = L.sum_(lmat).vec[0]
            dy = L.make_tensor(out.shape, data=np.ones(out.shape, np.float32))     # !sum backward: broadcast 1
            d = sce_bw(dy, z, y)
        else:  # mse-mean
            l = L.binary("sub", out, y)
            sq = L.arith_("mul", L.copy(l), l)
            loss = L.mean_(sq).vec[0]
            n = np.float32(sq.numel())
            dsq = L.make_tensor(out.shape, data=np.full(out.shape, np.float32(1.0) / n, np.float32))
            g1 = L.arith_("mul", L.copy(dsq), l)
            d = L.arith_("add", g1, L.arith_("mul", L.copy(dsq), l))
        gW, gb = [None] * nl, [None] * nl
        for i in range(nl - 1, -1, -1):
            xin = acts[i] if i == 0 else acts[i][0]
            dx, gW[i], gb[i] = linear_bw(d, xin, self.W[i])
            if i > 0:
                h_i, mask = acts[i]
                d = L.arith_("mul", dx, mask) if self.act == "relu" else sigmoid_bw(dx, h_i)
        return float(loss), gW, gb

    def apply(self, gW, gb, grad_scale=1.0):
        grads = [g.vec for g in gW] + [g.vec for g in gb]
        f = np.float32
        if grad_scale != 1.0:
            grads = [(g * f(grad_scale)).astype(np.float32) for g in grads]
        if self.opt == "sgd":
            for p, g in zip(self.params(), grads):
                p.vec[:] = p.vec - g * self.lr                    # (A-=B param (A*=scal grad lr))
            return
        self.n += 1
        n = self.n
        lr_t = self.lr * (np.sqrt(f(1.0) - self.b2 ** f(n)) / (f(1.0) - self.b1 ** f(n)))   # adam-step-lr
        lr_t = f(lr_t)
        for k, (p, g) in enumerate(zip(self.params(), grads)):
            m, v = self.m[k], self.v[k]
            m[:] = m + (g - m) * (f(1.0) - self.b1)               # Adam-Step-M
            v[:] = v + (f(1.0) - self.b2) * (g * g - v)           # Adam-Step-V
            p.vec[:] = p.vec - (m * lr_t) / (self.eps + np.sqrt(v))   # Adam-Step-Param

    def step(self, X, Y):
        loss, gW, gb = self.forward_backward(X, Y)
        self.apply(gW, gb)
        return loss
