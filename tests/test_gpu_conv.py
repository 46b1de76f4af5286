"""GPU parity of the unfold / Conv2D / pooling path through the C-ABI, against the oracle
(`oracle/conv.py`: the reference's own `waffe2_im2col_s_column` / `waffe2_col2im_s_column` / `waffe2_smax`
driven by a restatement of the Lisp composition).

Bars: Im2Col / Col2Im / unfold / fold / pooling forward+backward are data movement plus fixed-order fp32
sums -> bit-exact.  Conv2D goes through MatMulNode -> 1e-4 relative (the GEMM bar); bias gradients are
reductions -> 1e-5 of sum|x|."""
import ctypes as C
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _mods(gpu):
    return importlib.import_module("cl-waffe2_b200.nn"), importlib.import_module("cl-waffe2_b200.conv")


@pytest.fixture(params=[True, False], ids=["fused", "node-by-node"])
def fuse(request, gpu):
    nn, _ = _mods(gpu)
    old = (nn.FUSE, nn.CONCURRENT)
    nn.FUSE = nn.CONCURRENT = request.param
    yield request.param
    nn.FUSE, nn.CONCURRENT = old


# (N, C, Hin, Win, kh, kw, pad_h, pad_w, stride_h, stride_w)
GEOMS = [(2, 2, 2, 2, 2, 2, 0, 0, 1, 1),        # nn/t/conv.lisp:43-48
         (3, 3, 10, 10, 2, 2, 0, 0, 1, 1),      # nn/t/conv.lisp:5-7
         (1, 3, 16, 16, 4, 4, 0, 0, 4, 4),      # nn/t/conv.lisp:84-88 (pool window)
         (2, 3, 9, 7, 3, 2, 1, 1, 2, 1),        # padding + uneven strides: the node skips trailing rows
         (2, 2, 8, 8, 3, 3, 1, 2, 1, 2),
         (2, 2, 11, 9, 3, 3, 0, 0, 2, 3),
         (5, 4, 12, 12, 5, 5, 0, 0, 1, 1),      # MNIST-CNN second convolution
         (1, 1, 5, 5, 5, 5, 0, 0, 1, 1)]        # one window


@pytest.mark.parametrize("geo", GEOMS)
def test_im2col_col2im_nodes_bit_exact(gpu, geo):
    _, cv = _mods(gpu)
    from oracle import kernels
    K = kernels.load()
    N, Cc, Hin, Win, kh, kw, ph, pw, sh, sw = geo
    g = cv.Geometry((N, Cc, Hin, Win), (kh, kw), (sh, sw), (ph, pw))
    rng = np.random.default_rng(1)
    im = rng.standard_normal((N, Cc, g.H, g.W)).astype(np.float32)
    col0 = np.full((N, Cc, kh, kw, g.Ho, g.Wo), 7.0, np.float32)       # sentinel: untouched elements must stay
    want = col0.reshape(-1).copy()
    K.unfold("im2col", want, N, Cc, g.H, g.W, g.Ho, g.Wo, kh, kw, ph, pw, sh, sw, 1, 1, im.reshape(-1).copy())
    col = gpu.from_numpy(col0)
    cv.im2col(gpu.from_numpy(im), col, g)
    np.testing.assert_array_equal(col.to_numpy().reshape(-1), want)

    cin = rng.standard_normal(col0.shape).astype(np.float32)
    want_im = np.full(im.size, 3.0, np.float32)
    K.unfold("col2im", cin.reshape(-1).copy(), N, Cc, g.H, g.W, g.Ho, g.Wo, kh, kw, ph, pw, sh, sw, 1, 1, want_im)
    img = gpu.from_numpy(np.full(im.shape, 3.0, np.float32))
    cv.col2im(gpu.from_numpy(cin), img, g)
    np.testing.assert_array_equal(img.to_numpy().reshape(-1), want_im)


def test_unfold_golden_vectors(gpu, fuse):
    # nn/t/conv.lisp:43-62: unfold of ax+b (2 2 2 2) with a 2x2 kernel is 0..15; so is the gradient of
    # (!sum (!mul out (ax+b (2 8) 1 0)))
    nn, cv = _mods(gpu)
    img = gpu.ax_plus_b([2, 2, 2, 2], 1, 0)
    rows, g = cv.unfold(img, (1, 1), (2, 2), (1, 1), (0, 0))
    assert rows.shape == [2, 8]
    np.testing.assert_array_equal(rows.to_numpy().ravel(), np.arange(16, dtype=np.float32))
    dx = gpu.make_tensor([2, 2, 2, 2])
    cv._fold_into(gpu.ax_plus_b([2, 8], 1, 0), g, dx, False)
    np.testing.assert_array_equal(dx.to_numpy().ravel(), np.arange(16, dtype=np.float32))


@pytest.mark.parametrize("geo", GEOMS)
@pytest.mark.parametrize("layout", ["nchw", "channels-last"])
def test_unfold_fold_bit_exact(gpu, fuse, geo, layout):
    nn, cv = _mods(gpu)
    from oracle import loopnest as L, conv as OC
    N, Cc, Hin, Win, kh, kw, ph, pw, sh, sw = geo
    rng = np.random.default_rng(2)
    X = rng.standard_normal((N, Cc, Hin, Win)).astype(np.float32)
    ox = L.make_tensor(list(X.shape), data=X)
    orows, u = OC.unfold_fw(ox, (1, 1), (kh, kw), (sh, sw), (ph, pw))
    if layout == "nchw":
        x = gpu.from_numpy(X)
    else:
        x = cv._channels_last(N, Cc, Hin, Win)             # what Conv2D / MaxPool2D hand to the next layer
        importlib.import_module("cl-waffe2_b200.ops").move_(x, gpu.from_numpy(X))
    rows, g = cv.unfold(x, (1, 1), (kh, kw), (sh, sw), (ph, pw))
    np.testing.assert_array_equal(rows.to_numpy(), orows.to_numpy())
    D = rng.standard_normal(rows.shape).astype(np.float32)
    want = OC.unfold_bw(L.make_tensor(list(D.shape), data=D), u).to_numpy()
    dx = gpu.from_numpy(np.full(X.shape, 0.5, np.float32))
    cv._fold_into(gpu.from_numpy(D), g, dx, False)
    np.testing.assert_array_equal(dx.to_numpy(), want)
    cv._fold_into(gpu.from_numpy(D), g, dx, True)                      # gradient adder: dx += fold(D)
    np.testing.assert_array_equal(dx.to_numpy(), want + want)


@pytest.mark.parametrize("mode", ["max", "avg"])
@pytest.mark.parametrize("geo", [(1, 3, 16, 16, 4, 4, 0, 0, 4, 4), (4, 4, 26, 26, 2, 2, 0, 0, 2, 2), (2, 3, 9, 9, 3, 3, 0, 0, 2, 2),
                                 (2, 5, 8, 8, 2, 2, 1, 1, 2, 2), (3, 2, 7, 10, 3, 2, 0, 0, 1, 1), (2, 3, 12, 12, 5, 5, 0, 0, 1, 1)])
@pytest.mark.parametrize("layout", ["nchw", "channels-last"])
def test_pool2d_forward_backward_bit_exact(gpu, fuse, mode, geo, layout):
    nn, cv = _mods(gpu)
    from oracle import loopnest as L, conv as OC
    N, Cc, Hin, Win, kh, kw, ph, pw, sh, sw = geo
    rng = np.random.default_rng(3)
    X = rng.standard_normal((N, Cc, Hin, Win)).astype(np.float32)
    X[0, 0, :kh, :kw] = 1.5                                            # a tie inside one window: every tie gets the gradient
    ox = L.make_tensor(list(X.shape), data=X)
    ofw, obw = (OC.maxpool2d_fw, OC.maxpool2d_bw) if mode == "max" else (OC.avgpool2d_fw, OC.avgpool2d_bw)
    oout, octx = ofw(ox, (kh, kw), (sh, sw), (ph, pw))
    if layout == "nchw":
        x = gpu.from_numpy(X)
    else:                                              # what Conv2D hands over; its gradient has the same layout
        x = cv._channels_last(N, Cc, Hin, Win)
        importlib.import_module("cl-waffe2_b200.ops").move_(x, gpu.from_numpy(X))
        x._channels_last = True
        x._grad_like = cv._grad_like
    x.requires_grad = True
    tape = nn.Tape()
    nn._tape = tape
    try:
        out = (cv.maxpool2d if mode == "max" else cv.avgpool2d)(x, (kh, kw), (sh, sw), (ph, pw))
        assert out.shape == oout.shape
        if mode == "avg" and not fuse:
            # node by node `!mean` is the backend's tree reduction (1e-5 bar); the fused layer sums in window order
            np.testing.assert_allclose(out.to_numpy(), oout.to_numpy(), rtol=1e-5, atol=1e-6)
        else:
            np.testing.assert_array_equal(out.to_numpy(), oout.to_numpy())
        D = rng.standard_normal(out.shape).astype(np.float32)
        out.grad = cv._grad_like(out)
        importlib.import_module("cl-waffe2_b200.ops").move_(out.grad, gpu.from_numpy(D))
        out._has_grad = True
        for node in reversed(tape.nodes):
            node()
        nn._join_branches()
    finally:
        nn._tape = None
    want = obw(L.make_tensor(list(D.shape), data=D), octx).to_numpy()
    np.testing.assert_array_equal(x.grad.to_numpy(), want)


def test_pool2d_reference_gradient_constants(gpu, fuse):
    # nn/t/conv.lisp:84-110: (!mean (pool x)) on (1 3 16 16) with a 4x4 window: avg -> every gradient is
    # 0.0013020834; max -> 16 non-zero gradients per channel, each 0.020833334
    nn, cv = _mods(gpu)
    ops = importlib.import_module("cl-waffe2_b200.ops")
    rng = np.random.default_rng(4)
    for mode in ("avg", "max"):
        x = gpu.from_numpy(rng.standard_normal((1, 3, 16, 16)).astype(np.float32))
        x.requires_grad = True
        tape = nn.Tape()
        nn._tape = tape
        try:
            out = (cv.avgpool2d if mode == "avg" else cv.maxpool2d)(x, (4, 4))
            loss = nn.reduce_sum(out, mean=True)
            loss.grad = gpu.make_tensor(loss.shape)
            ops.fill_(loss.grad, 1.0)
            loss._has_grad = True
            for node in reversed(tape.nodes):
                node()
        finally:
            nn._tape = None
        g = x.grad.to_numpy()
        if mode == "avg":
            assert np.all(g == np.float32(0.0013020834))
        else:
            for k in range(3):
                nz = g[0, k][g[0, k] != 0]
                assert nz.size == 16 and np.all(nz == np.float32(0.020833334))


@pytest.mark.parametrize("geo", [(3, 3, 10, 10, 6, 2, 2, 0, 1), (4, 1, 28, 28, 4, 3, 3, 0, 1), (4, 4, 12, 12, 16, 5, 5, 0, 1),
                                 (2, 3, 9, 9, 5, 3, 3, 1, 2)])
def test_conv2d_forward_backward_matches_oracle(gpu, fuse, geo):
    nn, cv = _mods(gpu)
    ops = importlib.import_module("cl-waffe2_b200.ops")
    from oracle import loopnest as L, conv as OC
    N, Cin, H, W, Cout, kh, kw, pad, st = geo
    rng = np.random.default_rng(5)
    X = rng.standard_normal((N, Cin, H, W)).astype(np.float32)
    Wt, b = cv.conv_init_arrays(rng, Cin, Cout, (kh, kw))
    oout, octx = OC.conv2d_fw(L.make_tensor(list(X.shape), data=X), L.make_tensor(list(Wt.shape), data=Wt),
                              L.make_tensor([Cout], data=b), (st, st), (pad, pad))
    arena = nn.ParameterArena([Wt.shape, b.shape])
    arena.params[0].copy_from_numpy(Wt)
    arena.params[1].copy_from_numpy(b)
    x = gpu.from_numpy(X)
    x.requires_grad = True
    tape = nn.Tape()
    nn._tape = tape
    try:
        out = cv.conv2d(x, arena.params[0], arena.params[1], stride=st, padding_=pad)
        assert out.shape == oout.shape                           # nn/t/conv.lisp:5-7: (3 6 9 9)
        ref = oout.to_numpy()
        np.testing.assert_allclose(out.to_numpy(), ref, rtol=1e-4, atol=1e-4 * np.abs(ref).max())
        D = rng.standard_normal(out.shape).astype(np.float32)
        out.grad = cv._grad_like(out)
        ops.move_(out.grad, gpu.from_numpy(D))
        out._has_grad = True
        for node in reversed(tape.nodes):
            node()
        nn._join_branches()
        ops.sync()
    finally:
        nn._tape = None
    odx, odW, odb = OC.conv2d_bw(L.make_tensor(list(D.shape), data=D), octx)
    for got, want in ((x.grad.to_numpy(), odx.to_numpy()), (arena.params[0].grad.to_numpy(), odW.to_numpy())):
        np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-4 * np.abs(want).max())
    db = arena.params[1].grad.to_numpy().ravel()
    truth = D.astype(np.float64).sum(axis=(0, 2, 3))
    scale = np.abs(D).astype(np.float64).sum(axis=(0, 2, 3))
    assert np.all(np.abs(db - truth) <= 1e-5 * scale)
    assert np.all(np.abs(db - truth) <= np.abs(odb.to_numpy().ravel() - truth) + 1e-6 * scale)


def test_mnist_cnn_adam_matches_oracle(gpu, fuse):
    # examples/mnist/cnn.lisp:5-26 — the compiled (CUDA-graph) step against the oracle, per-step loss 1e-4
    nn, cv = _mods(gpu)
    from oracle.conv import OracleCNN
    rng = np.random.default_rng(0)
    B = 16
    X = rng.uniform(0, 1, size=(B, 1, 28, 28)).astype(np.float32)
    Y = np.eye(10, dtype=np.float32)[rng.integers(10, size=B)]
    init = cv.cnn_init_arrays(seed=3)
    model = cv.MNISTCNN(init=init)
    step = nn.CompiledStep(model, "softmax-cross-entropy-sum", nn.Adam(model.arena, lr=1e-3), X.shape, Y.shape)
    step.set_inputs(X, Y)
    ref = OracleCNN(init, lr=1e-3)
    ours, theirs = [], []
    for _ in range(8):
        step.step()
        ours.append(step.loss_value())
        theirs.append(ref.step(X, Y))
    assert step.graph is not None                 # replayed as a CUDA graph from the 2nd step on
    step.close()
    np.testing.assert_allclose(ours, theirs, rtol=1e-4)
    assert ours[-1] < ours[0]


def test_unfold_declines_dilation(gpu):
    _, cv = _mods(gpu)
    lib = gpu._lib.load()
    g = cv.Geometry((1, 1, 8, 8), (3, 3), (1, 1), (0, 0), (2, 2)).struct(False)
    x = gpu.make_tensor([1, 1, 8, 8])
    rows = gpu.make_tensor([64, 9])
    assert lib.wf2cu_unfold(C.byref(x.lower()), rows.vec.ptr, 0, C.byref(g)) == 3      # WF2CU_ERR_UNSUPPORTED


@pytest.mark.parametrize("shape", [(67600, 4, 9), (8100, 16, 100), (4096, 32, 128), (2048, 1, 7), (3001, 5, 33), (2500, 3, 1000), (5000, 20, 200)])
def test_skinny_forward_gemm(gpu, shape):
    # Conv2D's forward product (N*Ho*Wo, C*kh*kw) x (!t (C_out, C*kh*kw)) with bias + relu in the epilogue
    ops = importlib.import_module("cl-waffe2_b200.ops")
    m, n, k = shape
    rng = np.random.default_rng(7)
    A = rng.standard_normal((m, k)).astype(np.float32)
    B = rng.standard_normal((n, k)).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    ref = A.astype(np.float64) @ B.astype(np.float64).T
    out = ops.matmul(gpu.from_numpy(A), ops.t(gpu.from_numpy(B)))
    np.testing.assert_allclose(out.to_numpy(), ref, rtol=1e-4, atol=1e-4 * np.abs(ref).max())
    out = ops.matmul(gpu.from_numpy(A), ops.t(gpu.from_numpy(B)), bias=gpu.from_numpy(bias), relu=True)
    ref2 = np.maximum(ref + bias, 0)
    np.testing.assert_allclose(out.to_numpy(), ref2, rtol=1e-4, atol=1e-4 * np.abs(ref).max())


@pytest.mark.parametrize("shape", [(4, 9, 67600), (16, 100, 8100), (16, 128, 2048), (1, 1, 5000), (3, 17, 2049), (13, 77, 40000), (40, 50, 3000), (2, 170, 2500), (7, 36, 100000)])
def test_skinny_weight_gradient_gemm(gpu, shape):
    # Conv2D's weight gradient: (!t dout (K, m)) x rows (K, n), a long reduction with a small result
    ops = importlib.import_module("cl-waffe2_b200.ops")
    m, n, K = shape
    rng = np.random.default_rng(8)
    A = rng.standard_normal((K, m)).astype(np.float32)
    B = rng.standard_normal((K, n)).astype(np.float32)
    ref = A.astype(np.float64).T @ B.astype(np.float64)
    got = ops.matmul(ops.t(gpu.from_numpy(A)), gpu.from_numpy(B)).to_numpy()
    scale = np.abs(A).astype(np.float64).T @ np.abs(B).astype(np.float64)
    assert np.all(np.abs(got - ref) <= 1e-5 * scale)
    again = ops.matmul(ops.t(gpu.from_numpy(A)), gpu.from_numpy(B)).to_numpy()
    np.testing.assert_array_equal(got, again)                    # fixed fold order: run-to-run identical


@pytest.mark.parametrize("shape", [(67600, 4), (6400, 16), (1024, 10), (300, 64), (5000, 1), (777, 33)])
def test_narrow_column_sum(gpu, shape):
    # bias gradient of a layer with few outputs: sum over the rows of a tall narrow matrix
    ops = importlib.import_module("cl-waffe2_b200.ops")
    rows, cols = shape
    rng = np.random.default_rng(9)
    X = rng.standard_normal(shape).astype(np.float32)
    got = ops.sum(gpu.from_numpy(X), axis=0).to_numpy().ravel()
    truth = X.astype(np.float64).sum(axis=0)
    assert np.all(np.abs(got - truth) <= 1e-5 * np.abs(X).astype(np.float64).sum(axis=0))
    acc = gpu.from_numpy(np.ones((1, cols), np.float32))          # A+=B into existing values
    lib = gpu._lib.load()
    x = gpu.from_numpy(X)
    gb = gpu.view(acc, [":broadcast", rows], True)
    gpu._lib.check(lib.wf2cu_reduce_sum(C.byref(x.lower()), C.byref(gb.lower()), 1, 0.0))
    np.testing.assert_allclose(acc.to_numpy().ravel(), truth + 1.0, rtol=0, atol=1e-5 * np.abs(X).astype(np.float64).sum(axis=0).max())


@pytest.mark.parametrize("mode", [0, 1])
def test_pool2d_backward_applies_relu_mask(gpu, mode):
    # the `!relu` backward folded into the pooling backward: dx = fold(...) * (x > 0)
    _, cv = _mods(gpu)
    lib = gpu._lib.load()
    rng = np.random.default_rng(11)
    X = rng.standard_normal((3, 4, 10, 10)).astype(np.float32)
    g = cv.Geometry(X.shape, (2, 2), (2, 2), (0, 0))
    x = gpu.from_numpy(X)
    out = cv._channels_last(3, 4, 5, 5)
    gs = g.struct(False)
    gpu._lib.check(lib.wf2cu_pool2d(mode, C.byref(x.lower()), C.byref(out.lower()), C.byref(gs)))
    dout = gpu.from_numpy(rng.standard_normal((3, 4, 5, 5)).astype(np.float32))
    plain = gpu.make_tensor(X.shape)
    masked = gpu.from_numpy(np.ones(X.shape, np.float32))
    gpu._lib.check(lib.wf2cu_pool2d_backward(mode, C.byref(x.lower()), C.byref(out.lower()), C.byref(dout.lower()), C.byref(plain.lower()), C.byref(gs), 0, 0))
    gpu._lib.check(lib.wf2cu_pool2d_backward(mode, C.byref(x.lower()), C.byref(out.lower()), C.byref(dout.lower()), C.byref(masked.lower()), C.byref(gs), 1, 1))
    np.testing.assert_array_equal(masked.to_numpy(), np.float32(1.0) + plain.to_numpy() * (X > 0))
