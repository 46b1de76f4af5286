"""`:backward` compositions of the hot-path nodes (cl-waffe2_b200/backward.py).

CPU part: the compositions, run over the oracle's nodes (the reference's own C kernels), against the gradient goldens of
the reference's tests — `base-impl/t/arithmetic.lisp:25-51` (a = 10, b = 1: grads (1 1) (1 -1) (1 10) (1 -10)),
`base-impl/t/mathematical.lisp:36-141` (x = 1, dout = 1: abs' = signum, sign' = 0, sqrt' = 1/x, square' = x, sin' = cos,
cos' = -sin, tan' = 1/cos^2, sinh' = cosh, cosh' = -sinh [as the reference defines it], tanh' = 1/cosh^2, exp' = exp,
log2' = 1/(x ln 2), log10' = 1/(x ln 10), loge' = 1/x) — and against fp64 derivatives on random points.

GPU part (`-m gpu`): the same compositions on CUDATensor, every step one libwaffe2cuda.so kernel, against the oracle run:
bit-exact for the arithmetic nodes, within a few ulp for the chains that contain transcendentals.
(Run on a B200: profiles/r01d_pytest_gpu_backward_rules.log, 23 passed.)
"""
import importlib
import math

import numpy as np
import pytest

RNG = np.random.default_rng(5)


class OracleBackend:
    """cl-waffe2_b200.backward's backend protocol over oracle/loopnest.py."""

    def __init__(self, L):
        self.L = L

    def copy(self, x):
        return self.L.copy(x)

    def ones_like(self, x):
        return self.L.make_tensor(x.shape, data=np.ones(x.shape, np.float32))

    def arith_(self, op, x, y):
        if isinstance(y, (int, float)):
            return self.L.scalar_op_(op, x, float(y))
        return self.L.arith_(op, x, y)

    def unary(self, fn, x):
        # SignNode / SquareNode have LispTensor impls only (`backends/lisp/mathematics.lisp:49,53`: `signum`, `(* x x)`);
        # the C kernels name the natural logarithm `log`
        if fn == "sign":
            return self.L.make_tensor(x.shape, data=np.sign(x.to_numpy()).astype(np.float32))
        if fn == "square":
            v = x.to_numpy().astype(np.float32)
            return self.L.make_tensor(x.shape, data=v * v)
        return self.L.unary({"loge": "log"}.get(fn, fn), x)

    def expt(self, x, n):
        return self.L.expt(x, float(n))


@pytest.fixture(scope="module")
def bw(wf):
    return importlib.import_module("cl-waffe2_b200.backward")


def const(L, shape, v):
    return L.make_tensor(shape, data=np.full(shape, v, np.float32))


# ------------------------------------------------------------------ CPU: the reference's gradient goldens
@pytest.mark.parametrize("op,g1,g2", [("add", 1, 1), ("sub", 1, -1), ("mul", 1, 10), ("div", 1, -10)])
def test_arith_backward_goldens(bw, L, op, g1, g2):
    # define-arith-tester: a = 10, b = 1, dout = 1 => (grad a, grad b); note: mul's tester swaps because grad a = dout * b
    B = OracleBackend(L)
    a, b, dout = const(L, (100, 100), 10), const(L, (100, 100), 1), const(L, (100, 100), 1)
    ga, gb = bw.arith_backward(B, op, dout, a, b)
    assert (ga.to_numpy() == g1).all() and (gb.to_numpy() == g2).all()
    # inputs are untouched (every builder copies first, like !mul etc.)
    assert (a.to_numpy() == 10).all() and (b.to_numpy() == 1).all() and (dout.to_numpy() == 1).all()


GOLDEN_AT_ONE = {
    "abs": 1.0, "sign": 0.0, "sqrt": 1.0, "square": 1.0, "sin": math.cos(1), "cos": -math.sin(1),
    "tan": 1 / math.cos(1) ** 2, "sinh": math.cosh(1), "cosh": -math.sinh(1), "tanh": 1 / math.cosh(1) ** 2,
    "exp": math.exp(1), "log2": 1 / math.log(2), "log10": 1 / math.log(10), "loge": 1.0,
}


@pytest.mark.parametrize("fn", sorted(GOLDEN_AT_ONE))
def test_unary_backward_goldens_at_one(bw, L, fn):
    # define-mathematical-kernel-tester: (30 30) of ones, `~=` is a 1e-6-ish float comparison in the reference
    B = OracleBackend(L)
    x, dout = const(L, (30, 30), 1), const(L, (30, 30), 1)
    g = bw.unary_backward(B, fn, dout, x).to_numpy()
    assert g.shape == (30, 30)
    np.testing.assert_allclose(g, GOLDEN_AT_ONE[fn], rtol=2e-6, atol=1e-7)
    assert (x.to_numpy() == 1).all() and (dout.to_numpy() == 1).all()


TRUE_DERIVATIVE = {          # fp64; the four nodes whose reference formula is not the calculus one are excluded
    "abs": np.sign, "sin": np.cos, "cos": lambda x: -np.sin(x), "tan": lambda x: 1 / np.cos(x) ** 2,
    "sinh": np.cosh, "tanh": lambda x: 1 / np.cosh(x) ** 2, "asin": lambda x: 1 / np.sqrt(1 - x * x),
    "acos": lambda x: -1 / np.sqrt(1 - x * x), "atan": lambda x: 1 / (1 + x * x), "exp": np.exp,
    "log2": lambda x: 1 / (x * np.log(2)), "log10": lambda x: 1 / (x * np.log(10)), "loge": lambda x: 1 / x,
    "log1p": lambda x: 1 / (1 + x),
}


def _domain(fn, shape):
    if fn in ("log2", "log10", "loge", "sqrt"):
        return RNG.uniform(0.5, 3.0, size=shape).astype(np.float32)
    if fn == "log1p":
        return RNG.uniform(-0.5, 3.0, size=shape).astype(np.float32)
    if fn in ("asin", "acos"):
        return RNG.uniform(-0.9, 0.9, size=shape).astype(np.float32)
    return RNG.uniform(-1.0, 1.0, size=shape).astype(np.float32)


@pytest.mark.parametrize("fn", sorted(TRUE_DERIVATIVE))
def test_unary_backward_matches_fp64_derivative(bw, L, fn):
    B = OracleBackend(L)
    xs, ds = _domain(fn, (17, 33)), RNG.uniform(-2, 2, size=(17, 33)).astype(np.float32)
    g = bw.unary_backward(B, fn, L.make_tensor(xs.shape, data=ds), L.make_tensor(xs.shape, data=xs)).to_numpy()
    want = ds.astype(np.float64) * TRUE_DERIVATIVE[fn](xs.astype(np.float64))
    np.testing.assert_allclose(g, want, rtol=5e-6, atol=1e-6)


def test_reference_quirks_are_kept(bw, L):
    """cosh' = -sinh, sqrt' = 1/x, square' = x, sign' = x*0 (mathematics.lisp:131-147,184-188): reproduced as written."""
    B = OracleBackend(L)
    xs = RNG.uniform(0.5, 2.0, size=(8, 8)).astype(np.float32)
    x, one = L.make_tensor(xs.shape, data=xs), const(L, (8, 8), 1)
    np.testing.assert_allclose(bw.unary_backward(B, "cosh", one, x).to_numpy(), -np.sinh(xs.astype(np.float64)), rtol=2e-6)
    np.testing.assert_allclose(bw.unary_backward(B, "sqrt", one, x).to_numpy(), 1 / xs.astype(np.float64), rtol=2e-6)
    np.testing.assert_array_equal(bw.unary_backward(B, "square", one, x).to_numpy(), xs)
    assert (bw.unary_backward(B, "sign", one, x).to_numpy() == 0).all()


def test_expt_backward(bw, L):
    B = OracleBackend(L)
    xs = RNG.uniform(0.5, 2.0, size=(9, 11)).astype(np.float32)
    ds = RNG.uniform(-1, 1, size=(9, 11)).astype(np.float32)
    gx, gn = bw.expt_backward(B, L.make_tensor(xs.shape, data=ds), L.make_tensor(xs.shape, data=xs), 3.0, want_dn=True)
    x64, d64 = xs.astype(np.float64), ds.astype(np.float64)
    np.testing.assert_allclose(gx.to_numpy(), d64 * 3 * x64 ** 2, rtol=5e-6, atol=1e-6)
    np.testing.assert_allclose(gn.to_numpy(), d64 * x64 ** 3 * np.log(x64), rtol=5e-6, atol=1e-6)


def test_scalar_first_dispatch(bw, L):
    """(!sub 1 x) = (x * -1) + 1 and (!div 1 x) = (ones / x) * 1 (arithmetic.lisp:398-434, 179-208): same roundings as the
    reference's rewrite, not a fused 1 - x."""
    B = OracleBackend(L)
    xs = RNG.uniform(0.5, 2.0, size=(7, 5)).astype(np.float32)
    x = L.make_tensor(xs.shape, data=xs)
    np.testing.assert_array_equal(bw.sub(B, 1, x).to_numpy(), (xs * np.float32(-1)) + np.float32(1))
    np.testing.assert_array_equal(bw.div(B, 1, x).to_numpy(), (np.float32(1) / xs) * np.float32(1))
    np.testing.assert_array_equal(bw.add(B, 2, x).to_numpy(), xs + np.float32(2))
    np.testing.assert_array_equal(x.to_numpy(), xs)


# ------------------------------------------------------------------ CPU: the reference's chain-rule tests
def _sum_backward(B, dout_value, x):
    """`!sum` is `out*[:broadcast] += x` (reduction.lisp:60-79), so its gradient is dout broadcast over x's shape."""
    return B.arith_("mul", B.ones_like(x), float(dout_value))


def test_chain_rule_goldens(bw, L):
    """vm/generic-tensor/t/backward.lisp: chain-test3, 6, 7, 8, 9 and backward-being-not-destructed, replayed by applying the
    node :backward compositions by hand in reverse order (the VM does the same walk)."""
    B = OracleBackend(L)
    ones = const(L, (15, 15), 1)
    # chain-test3: (!div (!sum a) 2.0) -> grad a = 0.5 everywhere
    s = L.sum_(ones, axis=True)
    gs, _ = bw.arith_backward(B, "div", const(L, s.shape, 1), s, const(L, s.shape, 2))
    assert (_sum_backward(B, gs.to_numpy().ravel()[0], ones).to_numpy() == 0.5).all()
    # chain-test6: (!sum (!add (!sin a) b)) -> cos 1
    a = const(L, (15, 15), 1)
    g_add, _ = bw.arith_backward(B, "add", _sum_backward(B, 1.0, a), None, None)
    np.testing.assert_allclose(bw.unary_backward(B, "sin", g_add, a).to_numpy(), math.cos(1), rtol=2e-6)
    # chain-test7: (!sin (!sin a)) -> cos(sin 1) * cos 1
    x1 = B.unary("sin", a)
    g = bw.unary_backward(B, "sin", bw.unary_backward(B, "sin", const(L, (15, 15), 1), x1), a)
    np.testing.assert_allclose(g.to_numpy(), math.cos(math.sin(1)) * math.cos(1), rtol=2e-6)
    # chain-test8: (!sum (!add (!mul a b) c)), a = 3, b = 2, c = 1 -> 2, 3, 1
    a3, b2, c1 = const(L, (3, 3), 3), const(L, (3, 3), 2), const(L, (3, 3), 1)
    ab = bw.mul(B, a3, b2)
    d_ab, d_c = bw.arith_backward(B, "add", _sum_backward(B, 1.0, ab), ab, c1)
    d_a, d_b = bw.arith_backward(B, "mul", d_ab, a3, b2)
    assert (d_a.to_numpy() == 2).all() and (d_b.to_numpy() == 3).all() and (d_c.to_numpy() == 1).all()
    # chain-test9: prev = a*b + c ; (!sum (!add (!mul a1 prev) c1)), a1 = 2 -> a 4, b 6, c 2, a1 7, c1 1
    a1, c11 = const(L, (3, 3), 2), const(L, (3, 3), 1)
    prev = bw.add(B, ab, c1)
    top = bw.mul(B, a1, prev)
    d_top, d_c11 = bw.arith_backward(B, "add", _sum_backward(B, 1.0, top), top, c11)
    d_a1, d_prev = bw.arith_backward(B, "mul", d_top, a1, prev)
    d_ab2, d_c2 = bw.arith_backward(B, "add", d_prev, ab, c1)
    d_a2, d_b2 = bw.arith_backward(B, "mul", d_ab2, a3, b2)
    for got, want in [(d_a2, 4), (d_b2, 6), (d_c2, 2), (d_a1, 7), (d_c11, 1)]:
        assert (got.to_numpy() == want).all(), want
    # backward-being-not-destructed: parameters keep their values
    assert (a3.to_numpy() == 3).all() and (b2.to_numpy() == 2).all() and (c1.to_numpy() == 1).all() and (a1.to_numpy() == 2).all()


# ------------------------------------------------------------------ GPU: the same compositions on CUDATensor
@pytest.mark.gpu
@pytest.mark.parametrize("op", ["add", "sub", "mul", "div"])
def test_gpu_arith_backward_exact(gpu, bw, L, op):
    shape = (33, 130)
    a = RNG.uniform(-1, 1, size=shape).astype(np.float32)
    b = RNG.uniform(0.5, 2, size=shape).astype(np.float32)
    d = RNG.uniform(-1, 1, size=shape).astype(np.float32)
    mk_o = lambda v: L.make_tensor(shape, data=v)
    mk_g = lambda v: gpu.make_tensor(shape, data=v)
    oa, ob = bw.arith_backward(OracleBackend(L), op, mk_o(d), mk_o(a), mk_o(b))
    ga, gb = bw.arith_backward(bw.DeviceBackend(), op, mk_g(d), mk_g(a), mk_g(b))
    np.testing.assert_array_equal(ga.to_numpy(), oa.to_numpy())
    np.testing.assert_array_equal(gb.to_numpy(), ob.to_numpy())


@pytest.mark.gpu
@pytest.mark.parametrize("fn", ["abs", "sign", "sqrt", "square", "sin", "cos", "tan", "sinh", "cosh", "tanh", "asin", "acos", "atan",
                                "exp", "log2", "log10", "loge", "log1p"])
def test_gpu_unary_backward_matches_oracle(gpu, bw, L, fn):
    shape = (19, 67)
    xs, ds = _domain(fn, shape), RNG.uniform(-2, 2, size=shape).astype(np.float32)
    o = bw.unary_backward(OracleBackend(L), fn, L.make_tensor(shape, data=ds), L.make_tensor(shape, data=xs)).to_numpy()
    gx, gd = gpu.make_tensor(shape, data=xs), gpu.make_tensor(shape, data=ds)
    g = bw.unary_backward(bw.DeviceBackend(), fn, gd, gx).to_numpy()
    # each transcendental is within 2 ulp of the reference's; a chain of at most three of them plus exact arithmetic
    np.testing.assert_allclose(g, o, rtol=2e-6, atol=1e-7)
    np.testing.assert_array_equal(gx.to_numpy(), xs)        # saved inputs are not modified
    np.testing.assert_array_equal(gd.to_numpy(), ds)


@pytest.mark.gpu
def test_gpu_expt_backward_matches_oracle(gpu, bw, L):
    shape = (12, 40)
    xs = RNG.uniform(0.5, 2.0, size=shape).astype(np.float32)
    ds = RNG.uniform(-1, 1, size=shape).astype(np.float32)
    ox, on = bw.expt_backward(OracleBackend(L), L.make_tensor(shape, data=ds), L.make_tensor(shape, data=xs), 2.5, want_dn=True)
    gx, gn = bw.expt_backward(bw.DeviceBackend(), gpu.make_tensor(shape, data=ds), gpu.make_tensor(shape, data=xs), 2.5, want_dn=True)
    np.testing.assert_allclose(gx.to_numpy(), ox.to_numpy(), rtol=5e-6, atol=1e-7)
    np.testing.assert_allclose(gn.to_numpy(), on.to_numpy(), rtol=5e-6, atol=1e-7)


# ------------------------------------------------------------------ GPU vs the committed reference outputs
# tests/golden/ref_kernels_v1.npz holds what the reference's own compiled C kernels + SLEEF produced for fixed inputs
# (tests/golden/make_golden.py); these checks need neither /root/reference nor the compiled reference library.
import os

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_kernels_v1.npz")


def _ulps(a, b):
    ia = np.asarray(a, np.float32).ravel().view(np.int32).astype(np.int64)
    ib = np.asarray(b, np.float32).ravel().view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    return np.abs(ia - ib)


@pytest.mark.gpu
def test_gpu_matches_committed_reference_outputs(gpu):
    G = np.load(GOLDEN)
    two_d = lambda v: np.ascontiguousarray(v).reshape(5, -1)           # 525 = 5 x 105
    name_map = {"log": "loge"}
    for fn in ["sin", "cos", "tan", "asin", "acos", "atan", "sinh", "cosh", "tanh", "exp", "log", "log2", "log10", "log1p",
               "sqrt", "abs"]:
        x, want = two_d(G[f"unary/{fn}/x"]), G[f"unary/{fn}/y"]
        got = gpu.ops.unary(name_map.get(fn, fn), gpu.from_numpy(x)).to_numpy()
        worst = int(_ulps(got, want).max())
        assert worst <= 2, f"{fn}: {worst} ulp from the reference's output"
    for op in ["add", "sub", "mul", "div"]:                               # bit-exact
        gx = gpu.from_numpy(two_d(G["arith/x"]))
        getattr(gpu.ops, op + "_")(gx, gpu.from_numpy(two_d(G["arith/y"])))
        np.testing.assert_array_equal(gx.to_numpy().ravel(), G[f"arith/{op}"], err_msg=op)
    gx = gpu.from_numpy(two_d(G["pow/x"]))
    for p in (2.0, 3.0, 0.5, -1.5):
        worst = int(_ulps(gpu.ops.expt(gx, p).to_numpy(), G[f"pow/{p}"]).max())
        assert worst <= 2, f"pow {p}: {worst} ulp"
    cx, cy = gpu.from_numpy(two_d(G["cmp/x"])), gpu.from_numpy(two_d(G["cmp/y"]))
    for cmp in ["eq", "lt", "le", "gt", "ge"]:
        np.testing.assert_array_equal(gpu.ops.compare(cx, cmp, cy, 2.0, -1.0).to_numpy().ravel(), G[f"cmp/{cmp}"], err_msg=cmp)
        np.testing.assert_array_equal(gpu.ops.where(cx, cmp, 0.25, 1.0, 0.0).to_numpy().ravel(), G[f"cmp/{cmp}_scal"], err_msg=cmp)
    mx = gpu.from_numpy(G["maxmin/x"].reshape(1, -1))
    assert gpu.ops.max(mx).to_numpy().ravel()[0] == G["maxmin/max"][0]
    assert gpu.ops.min(mx).to_numpy().ravel()[0] == G["maxmin/min"][0]
    bx = gpu.from_numpy(two_d(G["bcast/x"]))
    gpu.ops.mul_(bx, float(G["bcast/y"][0]))
    np.testing.assert_array_equal(bx.to_numpy().ravel(), G["bcast/mul"])
    # `!sum`: the reference adds sequentially in fp32 (its result is the fixture); ours is a fixed tree.  Bar:
    # ours within 1e-5 * sum|x| of the fp64 truth, and the reference's own distance shown beside it.
    sx = G["seqsum/x"]
    truth, sumabs = float(sx.astype(np.float64).sum()), float(np.abs(sx.astype(np.float64)).sum())
    got = float(gpu.ops.sum(gpu.from_numpy(sx.reshape(1, -1)), axis=1).to_numpy().ravel()[0])
    ref = float(G["seqsum/sum"][0])
    assert abs(got - truth) <= 1e-5 * sumabs, (got, truth, ref)
    assert abs(got - truth) <= max(abs(ref - truth), 2e-7 * sumabs) * 4 + 1e-7 * sumabs, (got, truth, ref)
