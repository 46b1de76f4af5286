"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded
inputs.  Bars (BASELINE.json north_star): bit-exact for add/sub/mul/div/copy/compare and index
work; <= 2 ulp for transcendentals; 1e-5 relative (normalised by sum|x|) for reductions, judged
against an fp64 truth with the oracle's own error beside it; 1e-4 relative for GEMM.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RNG = np.random.default_rng(0)


def ulp_diff(a, b):
    a = np.asarray(a, np.float32).ravel()
    b = np.asarray(b, np.float32).ravel()
    ia = a.view(np.int32).astype(np.int64)
    ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    d = np.abs(ia - ib)
    both_nan = np.isnan(a) & np.isnan(b)
    d[both_nan] = 0
    return d


def pair(wf, L, shape, data=None, order="column"):
    """Same array on the device tensor and the oracle tensor."""
    if data is None:
        data = RNG.uniform(-1, 1, size=shape).astype(np.float32)
    return wf.make_tensor(shape, data=data, order=order), L.make_tensor(shape, data=data, order=order)


# ---------------------------------------------------------------- element-wise, exact
@pytest.mark.parametrize("op", ["add", "sub", "mul", "div"])
@pytest.mark.parametrize("shape", [(100, 100), (7, 13), (1, 1), (3, 5, 7), (33, 1030)])
def test_binary_inplace_exact(gpu, L, op, shape):
    gx, ox = pair(gpu, L, shape)
    gy, oy = pair(gpu, L, shape, RNG.uniform(0.5, 2, size=shape).astype(np.float32))
    getattr(gpu.ops, op + "_")(gx, gy)
    L.arith_(op, ox, oy)
    np.testing.assert_array_equal(gx.to_numpy(), ox.to_numpy())


@pytest.mark.parametrize("op", ["add", "mul"])
def test_binary_broadcast_views_exact(gpu, L, op):
    # !mul x (view y t (:broadcast n)) and the axis-0 twin — SURVEY §8d (ii),(iii)
    gx, ox = pair(gpu, L, (64, 48))
    gy, oy = pair(gpu, L, (64, 1))
    gz, oz = pair(gpu, L, (1, 48))
    r1 = getattr(gpu.ops, op)(gx, gpu.view(gy, True, [":broadcast", 48]))
    e1 = L.binary(op, ox, L.view(oy, True, [":broadcast", 48]))
    np.testing.assert_array_equal(r1.to_numpy(), e1.to_numpy())
    r2 = getattr(gpu.ops, op)(gx, gpu.view(gz, [":broadcast", 64], True))
    e2 = L.binary(op, ox, L.view(oz, [":broadcast", 64], True))
    np.testing.assert_array_equal(r2.to_numpy(), e2.to_numpy())


def test_sliced_and_stepped_views_exact(gpu, L):
    gx, ox = pair(gpu, L, (20, 36))
    gy, oy = pair(gpu, L, (20, 36))
    subs = [([2, 14], [4, 28]), ([0, 20, 2], True), (True, [0, 36, -1]), ([1, 19, -3], [2, 34, 4]), (3, [8, 16])]
    for s in subs:
        a, b = gpu.view(gx, *s), gpu.view(gy, *s)
        oa, ob = L.view(ox, *s), L.view(oy, *s)
        assert a.shape == oa.shape
        gpu.ops.add_(a, b)
        L.arith_("add", oa, ob)
    np.testing.assert_array_equal(gx.to_numpy(), ox.to_numpy())


def test_view_of_view_matches_oracle(gpu, L):
    gx = gpu.ax_plus_b((10, 12), 1, 0)
    ox = L.ax_plus_b((10, 12), 1, 0)
    g = gpu.view(gpu.view(gx, [1, 9], [2, 12]), [1, 5], [0, 6, 2])
    o = L.view(L.view(ox, [1, 9], [2, 12]), [1, 5], [0, 6, 2])
    assert g.shape == o.shape
    np.testing.assert_array_equal(g.to_numpy(), o.to_numpy())


def test_transposed_operand_exact(gpu, L):
    # !add x (!t y) — SURVEY §8d (viii): smem-staged transposed operand
    gx, ox = pair(gpu, L, (96, 160))
    gy, oy = pair(gpu, L, (160, 96))
    r = gpu.ops.add(gx, gpu.transpose(gy))
    e = L.binary("add", ox, L.transpose(oy))
    np.testing.assert_array_equal(r.to_numpy(), e.to_numpy())
    # permuted 4-D copy (GPT-2 head split): (2,3,8,5) -> swap axes 1,2
    g4, o4 = pair(gpu, L, (2, 3, 40, 33))
    gp = gpu.permute(g4, 3, 1, 2, 0)
    op_ = L.permute(o4, 3, 1, 2, 0)
    assert gp.shape == op_.shape
    np.testing.assert_array_equal(gpu.ops.copy(gp).to_numpy(), L.copy(op_).to_numpy())


def test_scalar_ops_exact(gpu, L):
    for op, s in [("add", 1.5), ("sub", 0.25), ("mul", -3.0), ("div", 7.0)]:
        gx, ox = pair(gpu, L, (17, 29))
        getattr(gpu.ops, op + "_")(gx, s)
        L.scalar_op_(op, ox, s)
        np.testing.assert_array_equal(gx.to_numpy(), ox.to_numpy())


def test_move_and_fill(gpu, L):
    gx, ox = pair(gpu, L, (9, 11))
    gy, oy = pair(gpu, L, (9, 11))
    gpu.ops.move_(gpu.view(gx, [0, 9, -1], True), gy)
    L.move_(L.view(ox, [0, 9, -1], True), oy)
    np.testing.assert_array_equal(gx.to_numpy(), ox.to_numpy())
    gpu.ops.fill_(gx, 2.5)
    assert (gx.to_numpy() == 2.5).all()


def test_nan_inf_zero_fill_semantics(gpu):
    # allocator zero-fills (LispTensor :initial-element 0); out*0 keeps NaN like the reference
    z = gpu.make_tensor((5, 5))
    assert (z.to_numpy() == 0).all()
    x = gpu.from_numpy(np.array([[np.nan, np.inf, 1.0]], np.float32))
    gpu.ops.mul_(x, 0.0)
    r = x.to_numpy()
    assert np.isnan(r[0, 0]) and np.isnan(r[0, 1]) and r[0, 2] == 0


# ---------------------------------------------------------------- compare
@pytest.mark.parametrize("cmp", ["eq", "lt", "le", "gt", "ge"])
def test_compare_exact(gpu, L, cmp):
    d = RNG.integers(-2, 3, size=(40, 24)).astype(np.float32)
    d2 = RNG.integers(-2, 3, size=(40, 24)).astype(np.float32)
    gx, ox = pair(gpu, L, (40, 24), d)
    gy, oy = pair(gpu, L, (40, 24), d2)
    np.testing.assert_array_equal(gpu.ops.where(gx, cmp, 0.0, 1.0, 0.0).to_numpy(), L.where_scal(cmp, ox, 0.0).to_numpy())
    np.testing.assert_array_equal(gpu.ops.compare(gx, cmp, gy, 2.0, -1.0).to_numpy(),
                                  L.compare(cmp, ox, oy, 2.0, -1.0).to_numpy())


# ---------------------------------------------------------------- unary math, <= 2 ulp
UNARY_DOMAINS = {
    "sin": (-20, 20), "cos": (-20, 20), "tan": (-1.5, 1.5), "asin": (-1, 1), "acos": (-1, 1), "atan": (-50, 50),
    "sinh": (-8, 8), "cosh": (-8, 8), "tanh": (-6, 6), "exp": (-30, 30), "log": (1e-6, 1e4),
    "log2": (1e-6, 1e4), "log10": (1e-6, 1e4), "log1p": (-0.9, 100), "sqrt": (0, 1e6), "abs": (-5, 5),
}
NAME_MAP = {"log": "loge"}


@pytest.mark.parametrize("fn", sorted(UNARY_DOMAINS))
def test_unary_within_2ulp(gpu, L, fn):
    lo, hi = UNARY_DOMAINS[fn]
    n = 1 << 16
    d = RNG.uniform(lo, hi, size=(n // 256, 256 + 3)).astype(np.float32)  # +3: exercises the oracle's n%8 tail
    gx, ox = pair(gpu, L, d.shape, d)
    r = gpu.ops.unary(NAME_MAP.get(fn, fn), gx).to_numpy()
    e = L.unary(fn, ox).to_numpy()
    worst = int(ulp_diff(r, e).max())
    assert worst <= 2, f"{fn}: {worst} ulp from the oracle"


def test_unary_special_values(gpu, L):
    d = np.array([[0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, 1e-38, 88.0, -104.0, 3.4e38, 1e-45, 0.5]], np.float32)
    for fn in ["sin", "cos", "exp", "tanh", "atan", "abs", "sqrt", "log"]:
        gx, ox = pair(gpu, L, d.shape, d)
        r = gpu.ops.unary(NAME_MAP.get(fn, fn), gx).to_numpy()
        with np.errstate(all="ignore"):
            e = L.unary(fn, ox).to_numpy()
        ok = (ulp_diff(r, e) <= 2) | (np.isnan(r.ravel()) & np.isnan(e.ravel()))
        assert ok.all(), f"{fn}: {r} vs {e}"


def test_expt(gpu, L):
    d = RNG.uniform(0.1, 4, size=(33, 40)).astype(np.float32)
    gx, ox = pair(gpu, L, d.shape, d)
    for p in (2.0, 3.0, 0.5, -1.5):
        worst = int(ulp_diff(gpu.ops.expt(gx, p).to_numpy(), L.expt(ox, p).to_numpy()).max())
        assert worst <= 2, f"pow {p}: {worst} ulp"


# ---------------------------------------------------------------- reductions
def _red_check(got, oracle, truth, sumabs, tol=1e-5):
    got = np.asarray(got, np.float64).ravel()
    oracle = np.asarray(oracle, np.float64).ravel()
    truth = np.asarray(truth, np.float64).ravel()
    sumabs = np.asarray(sumabs, np.float64).ravel() + 1e-30
    ours = (np.abs(got - truth) / sumabs).max()
    ref = (np.abs(oracle - truth) / sumabs).max()
    assert ours <= tol, f"ours-vs-fp64 {ours:.3e} (oracle-vs-fp64 {ref:.3e})"
    assert ours <= max(ref, 2e-7) * 4 + 1e-7, f"ours {ours:.3e} much worse than the reference's own {ref:.3e}"


@pytest.mark.parametrize("shape,axis", [((64, 300), 1), ((64, 300), 0), ((64, 300), True), ((5, 7, 9), 1),
                                        ((5, 7, 9), (0, 2)), ((4, 2048), 1), ((2000, 8), 0), ((3, 1031), -1),
                                        ((100, 100, 10), True), ((512, 4096), 1), ((300, 1000), 1), ((100, 8192), 1),
                                        ((70, 2048), -1), ((3, 100000), 1), ((1, 300000), True), ((4096, 512), 0),
                                        ((6, 64, 4096), 2), ((6, 300, 640), 1),
                                        # interleaved flat sum (ragged units), cluster/DSMEM column sums (ragged strips, batched)
                                        ((1, 1234567), True), ((2, 500000), 1), ((1024, 4096), 0), ((600, 2504), 0),
                                        ((3, 520, 1280), 1), ((777, 2500), 0), ((4096, 4096), 0), ((4096, 4096), 1)])
def test_sum_axes(gpu, L, shape, axis):
    d = RNG.normal(size=shape).astype(np.float32)
    gx, ox = pair(gpu, L, shape, d)
    r = gpu.ops.sum(gx, axis=axis)
    e = L.sum_(ox, axis=axis)
    assert r.shape == e.shape
    ax = None if axis is True else axis
    truth = d.astype(np.float64).sum(axis=ax, keepdims=True)
    sumabs = np.abs(d.astype(np.float64)).sum(axis=ax, keepdims=True)
    _red_check(r.to_numpy(), e.to_numpy(), truth, sumabs)


def test_sum_golden_ones(gpu):
    # base-impl/t/reduction.lisp:9-48: sums of ones over (100 100 100) -> exact counts
    x = gpu.make_tensor((100, 100, 100), initial_element=1.0)
    assert gpu.ops.sum(x).to_numpy().ravel()[0] == 1e6
    assert (gpu.ops.sum(x, axis=0).to_numpy() == 100).all()
    assert (gpu.ops.sum(x, axis=(1, 2)).to_numpy() == 1e4).all()
    assert (gpu.ops.mean(x, axis=1).to_numpy() == 1).all()


def test_sum_keepdims_and_inplace_accumulate(gpu, L):
    d = RNG.normal(size=(12, 40)).astype(np.float32)
    gx, ox = pair(gpu, L, d.shape, d)
    kb = gpu.ops.sum(gx, axis=1, keepdims=True)
    assert kb.shape == [12, 40]
    np.testing.assert_allclose(kb.to_numpy()[:, 0], d.sum(1), rtol=1e-5, atol=1e-5)
    # A+=B with a broadcast destination accumulates INTO the existing value
    go, oo = pair(gpu, L, (12, 1), np.arange(12, dtype=np.float32).reshape(12, 1))
    gpu.ops.add_(gpu.view(go, True, [":broadcast", 40]), gx)
    L.arith_("add", L.view(oo, True, [":broadcast", 40]), ox)
    np.testing.assert_allclose(go.to_numpy(), oo.to_numpy(), rtol=1e-5, atol=1e-5)
    # MulNode / SubNode / MoveTensorNode with a stride-0 destination
    for op in ("mul", "sub"):
        g1, o1 = pair(gpu, L, (12, 1), np.full((12, 1), 2.0, np.float32))
        dd = RNG.uniform(0.9, 1.1, size=(12, 40)).astype(np.float32)
        g2, o2 = pair(gpu, L, dd.shape, dd)
        getattr(gpu.ops, op + "_")(gpu.view(g1, True, [":broadcast", 40]), g2)
        L.arith_(op, L.view(o1, True, [":broadcast", 40]), o2)
        np.testing.assert_allclose(g1.to_numpy(), o1.to_numpy(), rtol=2e-5)
    g1, o1 = pair(gpu, L, (12, 1))
    gpu.ops.move_(gpu.view(g1, True, [":broadcast", 40]), gx)
    L.move_(L.view(o1, True, [":broadcast", 40]), ox)
    np.testing.assert_array_equal(g1.to_numpy(), o1.to_numpy())


@pytest.mark.parametrize("shape,axis", [((1024, 4096), 0), ((1024, 4096), 1), ((600, 2504), 0), ((1, 300000), 1), ((300, 1000), 1)])
def test_accumulating_reduction_large(gpu, L, shape, axis):
    """`A+=B` into a (:broadcast n) destination that already holds values (the reference's literal
    `!sum` path, reduction.lisp:60-79) on the shapes that take the tile / cluster / chunked kernels."""
    d = RNG.normal(size=shape).astype(np.float32)
    oshape = list(shape); oshape[axis] = 1
    init = RNG.normal(size=oshape).astype(np.float32)
    gx, ox = pair(gpu, L, shape, d)
    go, oo = pair(gpu, L, tuple(oshape), init)
    v = [True, True]; v[axis] = [":broadcast", shape[axis]]
    gpu.ops.add_(gpu.view(go, *v), gx)
    truth = init.astype(np.float64) + d.astype(np.float64).sum(axis=axis, keepdims=True)
    sumabs = np.abs(init.astype(np.float64)) + np.abs(d.astype(np.float64)).sum(axis=axis, keepdims=True)
    err = (np.abs(go.to_numpy().astype(np.float64) - truth) / sumabs).max()
    assert err <= 1e-5, err


def test_mean(gpu, L):
    d = RNG.uniform(0, 1, size=(50, 77)).astype(np.float32)
    gx, ox = pair(gpu, L, d.shape, d)
    for axis in (0, 1, True):
        r, e = gpu.ops.mean(gx, axis=axis), L.mean_(ox, axis=axis)
        np.testing.assert_allclose(r.to_numpy(), e.to_numpy(), rtol=1e-5)


# ---------------------------------------------------------------- softmax
@pytest.mark.parametrize("shape", [(16, 10), (8, 300), (6, 1024), (5, 2048), (3, 4096), (2, 9000), (7, 33)])
@pytest.mark.parametrize("avoid", [True, False])
def test_softmax_rows(gpu, L, shape, avoid):
    d = RNG.normal(size=shape).astype(np.float32)
    gx, ox = pair(gpu, L, shape, d)
    r = gpu.ops.softmax(gx, avoid_overflow=avoid).to_numpy()
    e = L.softmax(ox, avoid_overflow=avoid).to_numpy()
    d64 = d.astype(np.float64)
    sh = d64 - d64.mean(1, keepdims=True) if avoid else d64
    truth = np.exp(sh) / np.exp(sh).sum(1, keepdims=True)
    assert np.abs(r.sum(1) - 1).max() < 1e-4        # nn/t/activation.lisp:9-12
    ours = np.abs(r - truth).max() / truth.max()
    ref = np.abs(e - truth).max() / truth.max()
    assert ours <= 1e-5, (ours, ref)
    np.testing.assert_allclose(r, e, rtol=2e-5, atol=1e-9)


def test_softmax_axis_not_last(gpu, L):
    # GPT-2 quirk: (!softmax w) on a 4-D tensor normalises over axis 1 (heads)
    d = RNG.normal(size=(2, 12, 6, 10)).astype(np.float32)
    gx, ox = pair(gpu, L, d.shape, d)
    r = gpu.ops.softmax(gx, avoid_overflow=False, axis=1).to_numpy()
    e = L.softmax(ox, avoid_overflow=False, axis=1).to_numpy()
    np.testing.assert_allclose(r, e, rtol=2e-5, atol=1e-9)
    d2 = RNG.normal(size=(5, 40, 7)).astype(np.float32)
    gx, ox = pair(gpu, L, d2.shape, d2)
    np.testing.assert_allclose(gpu.ops.softmax(gx, axis=1).to_numpy(), L.softmax(ox, axis=1).to_numpy(), rtol=2e-5, atol=1e-9)


# ---------------------------------------------------------------- max / min / argmax
def test_maxmin_argmax(gpu, L):
    d = RNG.normal(size=(37, 50)).astype(np.float32)
    gx, ox = pair(gpu, L, d.shape, d)
    np.testing.assert_array_equal(gpu.ops.max(gx).to_numpy(), L.maxmin("max", ox).to_numpy())
    np.testing.assert_array_equal(gpu.ops.min(gx).to_numpy(), L.maxmin("min", ox).to_numpy())
    np.testing.assert_array_equal(gpu.ops.argmax(gx), L.argmaxmin("max", ox))
    np.testing.assert_array_equal(gpu.ops.argmin(gx), L.argmaxmin("min", ox))


# ---------------------------------------------------------------- fused composites
def test_fused_activations(gpu, L):
    d = RNG.normal(size=(64, 100)).astype(np.float32) * 3
    gx, ox = pair(gpu, L, d.shape, d)
    # sigmoid = (!div 1 (!add 1 (!exp (!mul -1 x))))
    e = L.scalar_op_("mul", L.copy(ox), -1.0)
    e = L.unary("exp", e)
    e = L.scalar_op_("add", e, 1.0)
    one = L.make_tensor(d.shape, data=np.ones(d.shape, np.float32))
    e = L.arith_("div", one, e)
    assert int(ulp_diff(gpu.ops.sigmoid(gx).to_numpy(), e.to_numpy()).max()) <= 4
    # relu = x * (x > 0)
    m = L.where_scal("gt", ox, 0.0)
    np.testing.assert_array_equal(gpu.ops.relu(gx).to_numpy(), L.arith_("mul", L.copy(ox), m).to_numpy())
    g = gpu.ops.gelu(gx).to_numpy()
    x64 = d.astype(np.float64)
    truth = 0.5 * x64 * (1 + np.tanh(np.sqrt(2 / np.pi) * (x64 + 0.044715 * x64 ** 3)))
    np.testing.assert_allclose(g, truth, rtol=1e-5, atol=1e-6)
    # the compiled gelu path (contiguous, 16-byte aligned) must equal the interpreted program on a strided view bit for bit
    d2 = RNG.normal(size=(64, 200)).astype(np.float32) * 3
    g2 = gpu.from_numpy(d2)
    half = gpu.view(g2, True, [0, 200, 2])                     # every other column: not vectorisable -> interpreter
    a = gpu.ops.gelu(gpu.from_numpy(np.ascontiguousarray(d2[:, ::2]))).to_numpy()
    b = gpu.ops.gelu(half).to_numpy()
    np.testing.assert_array_equal(a, b)
    # in place, as the GPT-2 example applies it
    g3 = gpu.from_numpy(d)
    gpu.ops.gelu(g3, out=g3)
    np.testing.assert_array_equal(g3.to_numpy(), g)


def test_softmax_over_heads_axis_vectorised(gpu, L):
    """GPT-2's `(!softmax w :avoid-overflow nil)` on a 4-D tensor normalises over axis 1 (examples/gpt-2/utils.lisp:236-242).
    Large enough for the 4-lanes-per-thread kernel; the scalar lane kernel (a strided view of the same data) is the check."""
    d = RNG.normal(size=(2, 12, 64, 2048)).astype(np.float32)
    gx = gpu.from_numpy(d)
    for avoid in (False, True):
        r = gpu.ops.softmax(gx, avoid_overflow=avoid, axis=1).to_numpy()
        x64 = d.astype(np.float64)
        if avoid:
            x64 = x64 - x64.mean(axis=1, keepdims=True)
        e = np.exp(x64)
        np.testing.assert_allclose(r, e / e.sum(axis=1, keepdims=True), rtol=2e-6, atol=1e-9)
        small = gpu.from_numpy(np.ascontiguousarray(d[:, :, :8, :64]))          # too small for the vector path
        np.testing.assert_array_equal(gpu.ops.softmax(small, avoid_overflow=avoid, axis=1).to_numpy(), r[:, :, :8, :64])
        # (!softmax (!div x n)) folded into one pass == the two nodes issued separately, bit for bit (both kernels)
        for t in (gx, small):
            two = gpu.ops.softmax(gpu.ops.div(t, 1024.0), avoid_overflow=avoid, axis=1).to_numpy()
            one = gpu.ops.softmax(t, avoid_overflow=avoid, axis=1, pre_divisor=1024.0).to_numpy()
            np.testing.assert_array_equal(one, two)
    # innermost axis: the library declines the folded form and the host issues the two nodes
    m = gpu.from_numpy(d[0, 0])
    np.testing.assert_array_equal(gpu.ops.softmax(m, axis=1, pre_divisor=7.0).to_numpy(),
                                  gpu.ops.softmax(gpu.ops.div(m, 7.0), axis=1).to_numpy())


# ---------------------------------------------------------------- matmul
def _sgemm_ref(a, b):
    return (a.astype(np.float32) @ b.astype(np.float32)).astype(np.float32)


def test_matmul_goldens(gpu):
    # base-impl/t/arithmetic.lisp:116-121: ax+b (3 3) @ itself
    a = gpu.ax_plus_b((3, 3), 1, 0)
    r = gpu.ops.matmul(a, a).to_numpy().ravel()
    np.testing.assert_array_equal(r, np.array([15, 18, 21, 42, 54, 66, 69, 90, 111], np.float32))


@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("m,n,k", [(3, 5, 4), (100, 10, 256), (130, 70, 33), (256, 256, 256)])
def test_matmul_transposes(gpu, ta, tb, m, n, k):
    A = RNG.uniform(-1, 1, size=(k, m) if ta else (m, k)).astype(np.float32)
    B = RNG.uniform(-1, 1, size=(n, k) if tb else (k, n)).astype(np.float32)
    ga, gb = gpu.from_numpy(A), gpu.from_numpy(B)
    r = gpu.ops.matmul(gpu.ops.t(ga) if ta else ga, gpu.ops.t(gb) if tb else gb).to_numpy()
    truth = (A.T if ta else A).astype(np.float64) @ (B.T if tb else B).astype(np.float64)
    err = np.abs(r - truth).max() / np.abs(truth).max()
    assert err <= 1e-4, err
    ref = np.abs(_sgemm_ref(A.T if ta else A, B.T if tb else B) - truth).max() / np.abs(truth).max()
    assert err <= max(4 * ref, 2e-6), (err, ref)


def test_matmul_batched(gpu):
    A = RNG.uniform(-1, 1, size=(2, 3, 20, 16)).astype(np.float32)
    B = RNG.uniform(-1, 1, size=(2, 3, 16, 24)).astype(np.float32)
    r = gpu.ops.matmul(gpu.from_numpy(A), gpu.from_numpy(B)).to_numpy()
    np.testing.assert_allclose(r, A @ B, rtol=1e-4, atol=1e-5)
    W = RNG.uniform(-1, 1, size=(24, 16)).astype(np.float32)
    r = gpu.ops.matmul(gpu.from_numpy(A), gpu.ops.t(gpu.from_numpy(W))).to_numpy()
    np.testing.assert_allclose(r, A @ W.T, rtol=1e-4, atol=1e-5)


@pytest.fixture
def tensor_gemm(gpu):
    """Force the tcgen05 3xTF32 path (mode 2) so small shapes cannot fall back to the SIMT kernel."""
    lib = gpu._lib.load()
    gpu._lib.check(lib.wf2cu_gemm_set_mode(2))
    yield gpu
    gpu._lib.check(lib.wf2cu_gemm_set_mode(0))


@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("m,n,k", [(128, 256, 32), (130, 70, 33), (100, 300, 77), (384, 520, 256), (1024, 256, 784)])
def test_matmul_tcgen05_3xtf32(tensor_gemm, ta, tb, m, n, k):
    """MatmulNode on the tensor cores vs fp64 truth; tolerance 1e-4 relative to max|C| (north_star),
    and within 2e-5 for k <= 2048 (measured model: 3e-7 + 8e-9 k)."""
    gpu = tensor_gemm
    A = RNG.uniform(-1, 1, size=(k, m) if ta else (m, k)).astype(np.float32)
    B = RNG.uniform(-1, 1, size=(n, k) if tb else (k, n)).astype(np.float32)
    ga, gb = gpu.from_numpy(A), gpu.from_numpy(B)
    r = gpu.ops.matmul(gpu.ops.t(ga) if ta else ga, gpu.ops.t(gb) if tb else gb).to_numpy()
    truth = (A.T if ta else A).astype(np.float64) @ (B.T if tb else B).astype(np.float64)
    err = np.abs(r - truth).max() / np.abs(truth).max()
    assert err <= 2e-5, err


@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("m,n,k", [(3, 5, 4), (128, 64, 32), (130, 70, 33), (100, 300, 77), (1024, 256, 784), (10, 256, 1024),
                                   (256, 784, 1024), (64, 64, 4096), (257, 129, 300)])
def test_matmul_tcgen05_fused_small_kernel(gpu, ta, tb, m, n, k):
    """mode 4: the one-launch tcgen05 kernel (in-kernel hi/lo split into swizzled shared memory, cluster
    split-K folded through DSMEM) on ragged and transposed shapes; bar 1e-4 (north_star), measured ~1e-6."""
    import ctypes as C
    lib = gpu._lib.load()
    gpu._lib.check(lib.wf2cu_gemm_set_mode(4))
    try:
        rng = np.random.default_rng(m * 7 + n * 3 + k)
        A = rng.uniform(-1, 1, size=((k, m) if ta else (m, k))).astype(np.float32)
        B = rng.uniform(-1, 1, size=((n, k) if tb else (k, n))).astype(np.float32)
        ga, gb = gpu.from_numpy(A), gpu.from_numpy(B)
        r = gpu.ops.matmul(gpu.ops.t(ga) if ta else ga, gpu.ops.t(gb) if tb else gb).to_numpy()
        truth = (A.T if ta else A).astype(np.float64) @ (B.T if tb else B).astype(np.float64)
        err = np.abs(r - truth).max() / np.abs(truth).max()
        assert err <= 2e-5, err
        # special values stay where they belong: an inf in A poisons exactly its row
        if m >= 3 and not ta:
            A2 = A.copy(); A2[1, 0] = np.inf
            r2 = gpu.ops.matmul(gpu.from_numpy(A2), gpu.ops.t(gb) if tb else gb).to_numpy()
            assert np.isfinite(r2[0]).all() and not np.isfinite(r2[1]).any()
    finally:
        gpu._lib.check(lib.wf2cu_gemm_set_mode(0))


def test_matmul_tcgen05_batched_and_shared_operand(tensor_gemm):
    gpu = tensor_gemm
    A = RNG.uniform(-1, 1, size=(2, 3, 200, 96)).astype(np.float32)
    B = RNG.uniform(-1, 1, size=(2, 3, 96, 264)).astype(np.float32)
    r = gpu.ops.matmul(gpu.from_numpy(A), gpu.from_numpy(B)).to_numpy()
    np.testing.assert_allclose(r, A.astype(np.float64) @ B.astype(np.float64), rtol=0, atol=2e-5 * 10)
    W = RNG.uniform(-1, 1, size=(520, 96)).astype(np.float32)      # LinearLayer: x @ (!t W), W shared by the batch
    r = gpu.ops.matmul(gpu.from_numpy(A), gpu.ops.t(gpu.from_numpy(W))).to_numpy()
    np.testing.assert_allclose(r, A.astype(np.float64) @ W.T.astype(np.float64), rtol=0, atol=2e-5 * 10)


def test_matmul_tcgen05_exact_small_integers(tensor_gemm):
    """The reference's golden style (base-impl/t/arithmetic.lisp:116-149): integer-valued operands give exact results."""
    gpu = tensor_gemm
    A = RNG.integers(-8, 9, size=(128, 64)).astype(np.float32)
    B = RNG.integers(-8, 9, size=(64, 256)).astype(np.float32)
    r = gpu.ops.matmul(gpu.from_numpy(A), gpu.from_numpy(B)).to_numpy()
    np.testing.assert_array_equal(r, A @ B)


def test_matmul_large_k_accumulation(tensor_gemm):
    gpu = tensor_gemm
    A = RNG.uniform(-1, 1, size=(128, 8192)).astype(np.float32)
    B = RNG.uniform(-1, 1, size=(8192, 256)).astype(np.float32)
    r = gpu.ops.matmul(gpu.from_numpy(A), gpu.from_numpy(B)).to_numpy()
    truth = A.astype(np.float64) @ B.astype(np.float64)
    assert np.abs(r - truth).max() / np.abs(truth).max() <= 1e-4


# ---------------------------------------------------------------- optimizers
def test_sgd_adam_steps(gpu):
    n = 1000
    p0 = RNG.normal(size=n).astype(np.float32)
    g0 = RNG.normal(size=n).astype(np.float32)
    lib = gpu._lib.load()
    p, g = gpu.from_numpy(p0), gpu.from_numpy(g0)
    gpu._lib.check(lib.wf2cu_sgd_step(p.vec.ptr, 0, g.vec.ptr, 0, n, 1e-3))
    np.testing.assert_array_equal(p.to_numpy(), (p0 - g0 * np.float32(1e-3)).astype(np.float32))
    p, m, v = gpu.from_numpy(p0), gpu.make_tensor((n,)), gpu.make_tensor((n,))
    b1, b2, eps, lr = np.float32(0.9), np.float32(0.999), np.float32(1e-7), np.float32(1e-3)
    gpu._lib.check(lib.wf2cu_adam_step(p.vec.ptr, 0, g.vec.ptr, 0, m.vec.ptr, 0, v.vec.ptr, 0, n, lr, None, b1, b2, eps, 1.0))
    one = np.float32(1)
    m0 = np.zeros(n, np.float32); v0 = np.zeros(n, np.float32)
    me = m0 + (g0 - m0) * (one - b1)                      # Adam-Step-M, adam.lisp:45-51
    ve = v0 + (one - b2) * (g0 * g0 - v0)                 # Adam-Step-V, :53-59
    pe = p0 - (me * lr) / (eps + np.sqrt(ve))             # Adam-Step-Param, :61-67
    np.testing.assert_array_equal(m.to_numpy(), me.astype(np.float32))
    np.testing.assert_array_equal(v.to_numpy(), ve.astype(np.float32))
    assert int(ulp_diff(p.to_numpy(), pe.astype(np.float32)).max()) <= 1
    # lr_t read from device memory (graph-replay path)
    p2, m2, v2 = gpu.from_numpy(p0), gpu.make_tensor((n,)), gpu.make_tensor((n,))
    lrd = gpu.from_numpy(np.array([lr], np.float32))
    gpu._lib.check(lib.wf2cu_adam_step(p2.vec.ptr, 0, g.vec.ptr, 0, m2.vec.ptr, 0, v2.vec.ptr, 0, n, 123.0, lrd.vec.ptr,
                                       b1, b2, eps, 1.0))
    np.testing.assert_array_equal(p2.to_numpy(), p.to_numpy())


# ---------------------------------------------------------------- errors, not fallbacks
def test_errors_are_loud(gpu):
    a, b = gpu.make_tensor((3, 4)), gpu.make_tensor((4, 3))
    with pytest.raises(ValueError):
        gpu.ops.add(a, b)
    lib = gpu._lib.load()
    va, vb = a.lower(), b.lower()
    import ctypes
    rc = lib.wf2cu_ew_binary(0, ctypes.byref(va), ctypes.byref(vb))
    assert rc == 2 and b"shape mismatch" in lib.wf2cu_last_error()
    with pytest.raises(gpu.Wf2CudaError):
        gpu._lib.check(lib.wf2cu_release(12345))


@pytest.mark.parametrize("ta,tb,m,n,k", [(0, 0, 512, 512, 256), (1, 0, 300, 700, 100), (0, 1, 1000, 517, 333), (1, 1, 257, 255, 48), (0, 0, 64, 64, 32)])
def test_matmul_tcgen05_256_tile_kernel(gpu, ta, tb, m, n, k):
    """modes 5 / 6: planes + the persistent 256x256 / 256x192-tile tcgen05 kernel (two M = 128 MMAs share each B slice,
    SWIZZLE_64B k-blocks of 16) on ragged and transposed shapes; bar 1e-4, and bit-identical to the 128x256-tile kernel
    (mode 2): all three add the same products in the same k order."""
    lib = gpu._lib.load()
    rng = np.random.default_rng(m + 3 * n + 5 * k)
    A = rng.uniform(-1, 1, size=((k, m) if ta else (m, k))).astype(np.float32)
    B = rng.uniform(-1, 1, size=((n, k) if tb else (k, n))).astype(np.float32)
    ga, gb = gpu.from_numpy(A), gpu.from_numpy(B)
    truth = (A.T if ta else A).astype(np.float64) @ (B.T if tb else B).astype(np.float64)
    outs = {}
    try:
        for mode in (5, 6, 2):
            gpu._lib.check(lib.wf2cu_gemm_set_mode(mode))
            outs[mode] = gpu.ops.matmul(gpu.ops.t(ga) if ta else ga, gpu.ops.t(gb) if tb else gb).to_numpy()
    finally:
        gpu._lib.check(lib.wf2cu_gemm_set_mode(0))
    assert np.abs(outs[5] - truth).max() <= 2e-5 * np.abs(truth).max()
    np.testing.assert_array_equal(outs[5], outs[2])
    np.testing.assert_array_equal(outs[6], outs[2])


@pytest.mark.parametrize("mode", [5, 6])
def test_matmul_tcgen05_256_tile_kernel_batched(gpu, mode):
    lib = gpu._lib.load()
    rng = np.random.default_rng(12)
    A = rng.uniform(-1, 1, size=(3, 300, 80)).astype(np.float32)
    W = rng.uniform(-1, 1, size=(520, 80)).astype(np.float32)        # shared by the batch, transposed
    gpu._lib.check(lib.wf2cu_gemm_set_mode(mode))
    try:
        r = gpu.ops.matmul(gpu.from_numpy(A), gpu.ops.t(gpu.from_numpy(W))).to_numpy()
    finally:
        gpu._lib.check(lib.wf2cu_gemm_set_mode(0))
    np.testing.assert_allclose(r, A.astype(np.float64) @ W.T.astype(np.float64), rtol=0, atol=2e-4)


@pytest.mark.parametrize("mode", [0, 2, 5, 6])
def test_matmul_tcgen05_persistent_kernels_fold_bias_and_relu(gpu, mode):
    """wf2cu_gemm_ex on the planes path: `+ bias` and `!relu` ride in the persistent kernels' TMEM -> global epilogue
    (the relu-backward mask stays a separate pass); equal to the unfused nodes bit for bit."""
    lib = gpu._lib.load()
    rng = np.random.default_rng(21)
    m, n, k = 1536, 1000, 1024                                  # 3.1 GFLOP: auto takes planes + persistent kernel
    A = rng.uniform(-1, 1, size=(m, k)).astype(np.float32)
    B = rng.uniform(-1, 1, size=(k, n)).astype(np.float32)
    bias = rng.uniform(-1, 1, size=(n,)).astype(np.float32)
    mask = rng.uniform(-1, 1, size=(m, n)).astype(np.float32)
    ga, gb, gbias, gmask = (gpu.from_numpy(x) for x in (A, B, bias, mask))
    gpu._lib.check(lib.wf2cu_gemm_set_mode(mode))
    try:
        plain = gpu.ops.matmul(ga, gb).to_numpy()
        fused = gpu.ops.matmul(ga, gb, bias=gbias, relu=True).to_numpy()
        masked = gpu.ops.matmul(ga, gb, bias=gbias, relu_mask=gmask).to_numpy()
    finally:
        gpu._lib.check(lib.wf2cu_gemm_set_mode(0))
    truth = A.astype(np.float64) @ B.astype(np.float64)
    assert np.abs(plain - truth).max() <= 2e-5 * np.abs(truth).max()
    want = plain + bias[None, :]
    np.testing.assert_array_equal(fused, want * (want > 0))
    np.testing.assert_array_equal(masked, want * (mask > 0))
