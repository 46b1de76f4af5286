"""Pins the oracle (oracle/loopnest.py + the reference's own C kernels in oracle/_ref) against the
golden vectors of the reference's test-suite.  CPU only.  Each test cites the reference test it
replays (paths under /root/reference/source)."""
import os
import numpy as np
import pytest

from oracle import kernels
from oracle import loopnest as L


# ---- vm/iterator/test-suites.lisp:14-166 -----------------------------------------------------
def test_range_size_goldens():
    assert L.range_size(L.make_range(0, 10, 1)) == 10
    assert L.range_size(L.make_range(0, 10, 2)) == 5
    assert L.range_size(L.make_range(0, 10, -2)) == 5


def _do_range(r):
    return [L.range_nth(r, i) for i in range(L.range_size(r))]


def test_do_range_goldens():
    assert _do_range(L.make_range(0, 10, 1)) == list(range(10))            # fixnum-do-range-test
    assert _do_range(L.make_range(0, 10, 2)) == [0, 2, 4, 6, 8]            # range-test-1
    assert _do_range(L.make_range(0, 10, -1)) == list(range(9, -1, -1))    # range-test-2 (pushed -> reversed)
    assert _do_range(L.make_range(2, 10, -2)) == [8, 6, 4, 2]              # range-test-3
    assert _do_range(L.make_range(1, 8, 1)) == list(range(1, 8))           # dynamic-do-range-test1
    assert _do_range(L.make_range(2, 10, 2)) == [2, 4, 6, 8]               # all-dynamic-do-range-test1
    assert _do_range(L.make_range(10, 2, 2)) == [8, 6, 4, 2]               # rev-dynamic-do-range-test


def test_range_compose_goldens():
    r = L.compose_range(L.make_range(2, 6, 2), L.make_range(2, 8, 1))       # compose-test
    assert (r.frm, r.to, r.step) == (4, 8, 2)
    r = L.compose_range(L.make_range(4), L.make_range(0, 5, -1))
    assert (r.frm, r.to, r.step) == (0, 1, 1)
    r = L.compose_range(L.make_range(0, 5), L.make_range(0, 5, -1))
    assert (r.frm, r.to, r.step) == (0, 5, 1)
    r = L.compose_range(L.make_range(0, 5, -1), L.make_range(0, 5))
    assert (r.frm, r.to, r.step) == (0, 5, -1)


@pytest.mark.parametrize("rng,expect", [((3, 0, -1), [0, 1, 2]), ((3, 0, 1), [2, 1, 0]), ((0, 3, 1), [0, 1, 2]),
                                        ((3, 6, 1), [3, 4, 5]), ((3, 6, -1), [5, 4, 3]), ((6, 3, 1), [5, 4, 3])])
def test_range_nth_goldens(rng, expect):
    assert [L.range_nth(L.make_range(*rng), i) for i in range(3)] == expect


# ---- vm/generic-tensor/t/view.lisp:30-70 -----------------------------------------------------
def _m(t, n):
    return [t.mref(i) for i in range(n)]


def test_view_goldens_1d():
    a = lambda: L.ax_plus_b([10], 1, 0)
    assert L.view(a(), 0).mref(0) == 0 and L.view(a(), 1).mref(0) == 1          # view-t->index-1d
    assert _m(L.view(a(), [0, 3]), 3) == [0, 1, 2]                               # view-t->slice-1d
    assert _m(L.view(a(), [1, 3]), 2) == [1, 2]
    assert _m(L.view(a(), [2, 10]), 8) == list(range(2, 10))
    assert _m(L.view(a(), [0, 10, 2]), 5) == [0, 2, 4, 6, 8]                     # view-t->slice-1d-by
    assert _m(L.view(a(), [2, -1, 2]), 4) == [2, 4, 6, 8]
    assert _m(L.view(a(), [2, 10, -2]), 4) == [8, 6, 4, 2]
    assert _m(L.view(L.view(a(), 1), 0), 1) == [1]                               # view-index->t-1d
    assert _m(L.view(L.view(a(), [1, 5]), [1, 3]), 2) == [2, 3]                  # view-slice->slice-1d
    assert _m(L.view(L.view(a(), [1, 5]), 1), 1) == [2]                          # view-slice->index-1d


def test_strides_column_is_c_order():
    # utils.lisp:50-67 + SURVEY §0 naming trap
    assert L.calc_strides([2, 3, 4], "column") == [12, 4, 1]
    assert L.calc_strides([2, 3, 4], "row") == [1, 2, 6]


# ---- base-impl/t/arithmetic.lisp:26-51,54-79 --------------------------------------------------
@pytest.mark.parametrize("op,result", [("add", 11), ("sub", 9), ("mul", 10), ("div", 10)])
def test_arith_testers(op, result):
    a = L.make_tensor([100, 100], data=np.full((100, 100), 10, np.float32))
    b = L.make_tensor([100, 100], data=np.ones((100, 100), np.float32))
    assert (L.binary(op, a, b).to_numpy() == result).all()
    assert (L.scalar_op_(op, L.copy(a), 1.0).to_numpy() == result).all()         # scalar-mat testers


def test_move_tester():
    a = L.make_tensor([100, 100], data=np.full((100, 100), 10, np.float32))
    b = L.make_tensor([100, 100], data=np.ones((100, 100), np.float32))
    assert (L.move_(a, b).to_numpy() == 1).all()


# ---- base-impl/t/arithmetic.lisp:116-149 (matmul goldens) -------------------------------------
def test_matmul_goldens():
    a, b = L.ax_plus_b([3, 3], 1, 0), L.ax_plus_b([3, 3], 1, 0)
    assert L.matmul(a, b).vec.tolist() == [15, 18, 21, 42, 54, 66, 69, 90, 111]
    assert L.matmul(a, b, ta=True, tb=True).vec.tolist() == [15, 42, 69, 18, 54, 90, 21, 66, 111]
    a, b = L.ax_plus_b([3, 4], 1, 0), L.ax_plus_b([3, 4], 1, 0)
    assert L.matmul(a, b, tb=True).vec.tolist() == [14, 38, 62, 38, 126, 214, 62, 214, 366]
    assert L.matmul(a, b, ta=True).vec.tolist() == [80, 92, 104, 116, 92, 107, 122, 137, 104, 122, 140, 158,
                                                    116, 137, 158, 179]


# ---- base-impl/t/arithmetic.lisp:236-357 (max/min, compares) ----------------------------------
def test_maxmin_goldens():
    a = L.ax_plus_b([9, 9], 1, 0)
    assert L.maxmin("max", a).vec.tolist() == [8, 17, 26, 35, 44, 53, 62, 71, 80]
    assert L.maxmin("min", a).vec.tolist() == [0, 9, 18, 27, 36, 45, 54, 63, 72]


@pytest.mark.parametrize("cmp,av,bv,expect", [("lt", 1, 2, 1), ("lt", 2, 1, 0), ("le", 1, 1, 1), ("le", 2, 1, 0),
                                              ("gt", 1, 2, 0), ("gt", 2, 1, 1), ("ge", 1, 1, 1), ("ge", 2, 1, 1),
                                              ("eq", 1, 1, 1)])
def test_compare_goldens(cmp, av, bv, expect):
    a, b = L.ax_plus_b([9, 9], 0, av), L.ax_plus_b([9, 9], 0, bv)
    assert (L.compare(cmp, a, b).to_numpy() == expect).all()
    assert (L.where_scal(cmp, a, float(bv)).to_numpy() == expect).all()


# ---- base-impl/t/reduction.lisp:9-48 -----------------------------------------------------------
def test_sum_mean_goldens():
    a = L.make_tensor([100, 100, 100], data=np.ones((100, 100, 100), np.float32))
    assert L.sum_(a).mref(0, 0, 0) == 1e6
    assert L.sum_(a, axis=0).mref(0, 0, 0) == 100
    assert L.sum_(a, axis=1).mref(0, 0, 0) == 100
    assert L.sum_(a, axis=-1).mref(0, 0, 0) == 100
    assert L.sum_(a, axis=[0, 1]).mref(0, 0, 0) == 10000
    for ax in (True, 0, 1, -1, [0, 1]):
        assert L.mean_(a, axis=ax).mref(0, 0, 0) == 1.0


def test_sum_is_sequential_fp32():
    # cl-waffe2-simd.c:159,205-212: stride-0 destination => strictly sequential accumulation
    x = np.random.default_rng(1).normal(size=(1, 5000)).astype(np.float32)
    got = L.sum_(L.make_tensor([1, 5000], data=x)).vec[0]
    seq = np.float32(0)
    for v in x.ravel():
        seq = np.float32(seq + v)
    assert got == seq


# ---- base-impl/t/mathematical.lisp:37-142 (f(1) with the suite's ~= tolerance) -----------------
@pytest.mark.parametrize("fn,ref", [("abs", abs), ("sin", np.sin), ("cos", np.cos), ("tan", np.tan), ("sinh", np.sinh),
                                    ("cosh", np.cosh), ("tanh", np.tanh), ("exp", np.exp), ("log2", np.log2),
                                    ("log10", np.log10), ("log", np.log), ("sqrt", np.sqrt)])
def test_math_at_one(fn, ref):
    a = L.make_tensor([30, 30], data=np.ones((30, 30), np.float32))
    out = L.unary(fn, a).to_numpy()
    assert np.all(np.abs(out - ref(1.0)) < 1e-5)


# ---- nn/t/activation.lisp:4-19 -----------------------------------------------------------------
def test_softmax_rows_sum_to_one():
    x = np.random.default_rng(2).normal(size=(10, 10)).astype(np.float32)
    s = L.softmax(L.make_tensor([10, 10], data=x)).to_numpy()
    assert np.all(np.abs(s.sum(1) - 1) < 1e-4)


# ---- loop-plan shape: contiguous collapses to ONE call, broadcast does not ------
def test_loop_plan_call_counts():
    x = L.make_tensor([64, 32]); y = L.make_tensor([64, 32])
    st = {}
    L.arith_("add", x, y, stats=st)
    assert st["leaf_calls"] == 1
    col = L.view(L.make_tensor([64, 1]), True, [":broadcast", 32])
    st = {}
    L.arith_("mul", x, col, stats=st)
    assert st["leaf_calls"] == 64
    st = {}
    L.sum_(x, axis=True, stats=st)
    assert st["leaf_calls"] == 1      # (:broadcast 64) x (:broadcast 32) destination collapses too


# ---- port vs the reference build ------------------------------------------------------------------
@pytest.mark.skipif(not kernels.have_reference(), reason="oracle/_ref not built (needs /root/reference)")
def test_port_matches_reference_build():
    ref, port = kernels.load("reference"), kernels.load("port")
    assert ref.kind == "reference" and port.kind == "port"
    rng = np.random.default_rng(3)
    n = 1003
    for name in kernels.ARITH:
        x1 = rng.uniform(0.5, 2, n).astype(np.float32); y = rng.uniform(0.5, 2, n).astype(np.float32)
        x2 = x1.copy()
        ref.arith(name, n, x1, 0, 1, y, 0, 1); port.arith(name, n, x2, 0, 1, y, 0, 1)
        assert np.array_equal(x1, x2)
    for name in kernels.UNARY:
        x = rng.uniform(0.1, 0.9, n).astype(np.float32)
        o1 = np.zeros(n, np.float32); o2 = np.zeros(n, np.float32)
        ref.unary(name, n, x, 0, 1, o1, 0, 1); port.unary(name, n, x, 0, 1, o2, 0, 1)
        d = np.abs(o1.view(np.int32).astype(np.int64) - o2.view(np.int32).astype(np.int64))
        assert d.max() <= 1, (name, d.max())      # SLEEF u10 body vs double-libm tail arithmetic
        assert np.array_equal(o1[-(n % 8):], o2[-(n % 8):])   # the reference's own tail IS the port's arithmetic


# ---- committed fixtures: outputs of the reference's own compiled kernels (tests/golden/make_golden.py) ------------------
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_kernels_v1.npz")


def _ulps(a, b):
    ia = a.view(np.int32).astype(np.int64)
    ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    return np.abs(ia - ib)


@pytest.mark.parametrize("which", ["port", "reference"])
def test_kernels_reproduce_committed_golden_outputs(which):
    """`port`: oracle/port.c against what the reference's C kernels + SLEEF produced (bit-exact for everything except the
    transcendentals' SIMD body, where SLEEF's polynomial and the port's double-libm differ by <= 1 ulp).
    `reference`: the compiled reference itself must reproduce the fixture bit for bit (a stale fixture fails here)."""
    if which == "reference" and not kernels.have_reference():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    K = kernels.load(which)
    assert K.kind == which
    G = np.load(GOLDEN)
    exact = which == "reference"
    for fn in ["sin", "cos", "tan", "asin", "acos", "atan", "sinh", "cosh", "tanh", "exp", "log", "log2", "log10", "log1p",
               "sqrt", "abs"]:
        x, want = G[f"unary/{fn}/x"].copy(), G[f"unary/{fn}/y"]
        got = np.zeros_like(want)
        K.unary(fn, len(x), x, 0, 1, got, 0, 1)
        assert _ulps(got, want).max() <= (0 if exact else 1), (fn, _ulps(got, want).max())
    x, y = G["arith/x"], G["arith/y"].copy()
    for op in kernels.ARITH:
        acc = x.copy()
        K.arith(op, len(acc), acc, 0, 1, y, 0, 1)
        assert np.array_equal(acc, G[f"arith/{op}"]), op
    px = G["pow/x"].copy()
    for p in (2.0, 3.0, 0.5, -1.5):
        got = np.zeros_like(px)
        K.pow(len(px), px, 0, 1, p, got, 0, 1)
        assert _ulps(got, G[f"pow/{p}"]).max() <= (0 if exact else 1), p
    cx, cy = G["cmp/x"].copy(), G["cmp/y"].copy()
    for op in kernels.CMP:
        got = np.zeros_like(cx)
        K.cmp(op, len(cx), cx, 0, 1, cy, 0, 1, got, 0, 1, 2.0, -1.0)
        assert np.array_equal(got, G[f"cmp/{op}"]), op
        got = np.zeros_like(cx)
        K.cmp_scal(op, len(cx), cx, 0, 1, got, 0, 1, 0.25, 1.0, 0.0)
        assert np.array_equal(got, G[f"cmp/{op}_scal"]), op
    mx = G["maxmin/x"].copy()
    for name in ("max", "min"):
        got = np.zeros(1, np.float32)
        K.maxmin(name, len(mx), mx, 0, 1, got, 0)
        assert got[0] == G[f"maxmin/{name}"][0]
    sx = G["seqsum/x"].copy()
    acc = np.zeros(1, np.float32)
    K.arith("add", len(sx), acc, 0, 0, sx, 0, 1)
    assert acc[0] == G["seqsum/sum"][0]                       # the sequential fp32 order of `!sum`
    seq = np.float32(0)
    for v in sx:
        seq = np.float32(seq + v)
    assert seq == G["seqsum/sum"][0]
    bx = G["bcast/x"].copy()
    K.arith("mul", len(bx), bx, 0, 1, G["bcast/y"].copy(), 0, 0)
    assert np.array_equal(bx, G["bcast/mul"])
    geo = [int(v) for v in G["unfold/geometry"]]
    col = np.zeros_like(G["unfold/col"])
    K.unfold("im2col", col, *geo, G["unfold/img"].copy())
    assert np.array_equal(col, G["unfold/col"])
    back = np.zeros_like(G["unfold/back"])
    K.unfold("col2im", G["unfold/col"].copy(), *geo, back)
    assert np.array_equal(back, G["unfold/back"])


# ---- nn/t/regression.lisp:429-454 (grad-decay-test) and backends/lisp/t/package.lisp:26-32 (lazy-test) ---------------------
@pytest.mark.parametrize("opt,lr", [("sgd", 1e-3), ("adam", 1e-3)])
def test_mlp_loss_decreases_on_sin_wave(opt, lr):
    """X = sin(ax+b (100 100) 0.01 0.1), Y = cos(ax+b (100 1) 0.01 0.1), 100 -> 10 -> 1 sigmoid MLP, MSE: the loss after
    training is below the first one (the reference runs 3000 SGD steps; 300 already show it)."""
    import importlib
    from oracle.mlp import OracleMLP
    nn = importlib.import_module("cl-waffe2_b200.nn")
    X = L.unary("sin", L.ax_plus_b([100, 100], 0.01, 0.1)).to_numpy()
    Y = L.unary("cos", L.ax_plus_b([100, 1], 0.01, 0.1)).to_numpy()
    m = OracleMLP([100, 10, 1], "sigmoid", nn.mlp_init_arrays([100, 10, 1], seed=1), "mse-mean", opt=opt, lr=lr)
    losses = [m.step(X, Y) for _ in range(300)]
    assert np.isfinite(losses).all()
    assert losses[-1] < losses[0]
    assert losses[-1] < losses[len(losses) // 2] < losses[0] or opt == "adam"      # monotone trend for plain SGD


def test_sin_of_permuted_tensor_matches_elementwise_truth():
    """lazy-test: (!sin (!permute (ax+b (10 10 10) 1 0) 2 0 1)) through the CPU kernels equals sin applied element by element
    (the LispTensor route) within the test's 1e-5."""
    t = L.permute(L.ax_plus_b([10, 10, 10], 1, 0), 2, 0, 1)
    got = L.unary("sin", t).to_numpy()
    want = np.sin(t.to_numpy().astype(np.float64))
    assert got.shape == want.shape == tuple(t.shape)
    assert np.abs(got - want).max() < 1e-5
