"""CPU: pin the unfold / Conv2D / pooling oracle (`oracle/conv.py`) against the reference's own golden
vectors (`/root/reference/source/nn/t/conv.lisp`) and the scalar port against the reference's C build."""
import math

import numpy as np
import pytest

from oracle import conv as OC
from oracle import kernels
from oracle import loopnest as L


def test_unfold_forward_golden():
    # nn/t/conv.lisp:43-48
    img = L.ax_plus_b([2, 2, 2, 2], 1, 0)
    rows, _ = OC.unfold_fw(img, (1, 1), (2, 2), (1, 1), (0, 0))
    assert rows.shape == [2, 8]
    assert np.array_equal(rows.to_numpy().ravel(), np.arange(16, dtype=np.float32))


def test_unfold_backward_golden():
    # nn/t/conv.lisp:50-62: d/dimg (!sum (!mul out (ax+b (2 8) 1 0))) = 0..15
    img = L.ax_plus_b([2, 2, 2, 2], 1, 0)
    _, u = OC.unfold_fw(img, (1, 1), (2, 2), (1, 1), (0, 0))
    dx = OC.unfold_bw(L.ax_plus_b([2, 8], 1, 0), u)
    assert np.array_equal(dx.to_numpy().ravel(), np.arange(16, dtype=np.float32))


def test_conv2d_shape_and_gradients():
    # nn/t/conv.lisp:5-7 (shape (3 6 9 9)), :9-33 (weight gradients of one window are not all equal)
    rng = np.random.default_rng(0)
    X = (np.float32(0.001) * np.arange(3 * 3 * 10 * 10, dtype=np.float32)).reshape(3, 3, 10, 10)
    a = math.sqrt(1 / 12)
    W = rng.uniform(-a, a, size=(6, 3, 2, 2)).astype(np.float32)
    b = rng.uniform(-a, a, size=(6,)).astype(np.float32)
    out, ctx = OC.conv2d_fw(L.make_tensor([3, 3, 10, 10], data=X), L.make_tensor([6, 3, 2, 2], data=W), L.make_tensor([6], data=b))
    assert out.shape == [3, 6, 9, 9]
    n = out.numel()
    dout = L.make_tensor(out.shape, data=np.full(out.shape, np.float32(1) / np.float32(n), np.float32))   # !mean backward
    dx, dW, db = OC.conv2d_bw(dout, ctx)
    g = dW.to_numpy()
    assert any(not np.all(g[i, k] == g[i, k].ravel()[0]) for i in range(6) for k in range(3))
    # cross-check the whole layer against a direct fp64 convolution
    ref = np.zeros((3, 6, 9, 9))
    for i in range(2):
        for j in range(2):
            ref += np.einsum("nchw,oc->nohw", X[:, :, i:i + 9, j:j + 9].astype(np.float64), W[:, :, i, j].astype(np.float64))
    ref += b.reshape(1, 6, 1, 1)
    assert np.abs(out.to_numpy() - ref).max() < 1e-5
    assert np.allclose(db.to_numpy().ravel(), np.full(6, 3 * 81 / n), rtol=1e-5)


@pytest.mark.parametrize("avg", [True, False])
def test_pool2d_gradient_constants(avg):
    # nn/t/conv.lisp:84-110
    rng = np.random.default_rng(1)
    x = L.make_tensor([1, 3, 16, 16], data=rng.standard_normal((1, 3, 16, 16)).astype(np.float32))
    fw, bw = (OC.avgpool2d_fw, OC.avgpool2d_bw) if avg else (OC.maxpool2d_fw, OC.maxpool2d_bw)
    out, ctx = fw(x, (4, 4))
    assert out.shape == [1, 3, 4, 4]
    n = out.numel()
    dout = L.make_tensor(out.shape, data=np.full(out.shape, np.float32(1) / np.float32(n), np.float32))
    g = bw(dout, ctx).to_numpy()
    if avg:
        assert np.all(g == np.float32(0.0013020834))
    else:
        for k in range(3):
            nz = g[0, k][g[0, k] != 0]
            assert nz.size == 16 and np.all(nz == np.float32(0.020833334))
        assert np.array_equal(out.to_numpy(), x.to_numpy().reshape(1, 3, 4, 4, 4, 4).max(axis=(3, 5)))


@pytest.mark.skipif(not kernels.have_reference(), reason="oracle/_ref not built (needs /root/reference)")
def test_unfold_port_matches_reference_build():
    ref, port = kernels.load("reference"), kernels.load("port")
    rng = np.random.default_rng(0)
    for (N, C, Hin, Win, KH, KW, ph, pw, sh, sw) in [(2, 2, 2, 2, 2, 2, 0, 0, 1, 1), (3, 3, 10, 10, 2, 2, 0, 0, 1, 1),
                                                     (2, 3, 9, 7, 3, 2, 1, 1, 2, 1), (1, 3, 16, 16, 4, 4, 0, 0, 4, 4),
                                                     (2, 2, 8, 8, 3, 3, 1, 2, 1, 2), (2, 2, 11, 9, 3, 3, 0, 0, 2, 3)]:
        Ho = OC.conv_out_size(Hin, ph, 1, KH, sh)
        Wo = OC.conv_out_size(Win, pw, 1, KW, sw)
        H, W = Hin + 2 * ph + Ho % sh, Win + 2 * pw + Wo % sw
        im = rng.standard_normal(N * C * H * W).astype(np.float32)
        c1 = np.full(N * C * KH * KW * Ho * Wo, 7.0, np.float32)
        c2 = c1.copy()
        ref.unfold("im2col", c1, N, C, H, W, Ho, Wo, KH, KW, ph, pw, sh, sw, 1, 1, im)
        port.unfold("im2col", c2, N, C, H, W, Ho, Wo, KH, KW, ph, pw, sh, sw, 1, 1, im)
        assert np.array_equal(c1, c2)
        if ph or pw:
            assert (c1 == 7.0).any()          # the reference's node skips trailing rows/columns when pad > 0
        col = rng.standard_normal(c1.size).astype(np.float32)
        i1 = np.full(im.size, 3.0, np.float32)
        i2 = i1.copy()
        ref.unfold("col2im", col, N, C, H, W, Ho, Wo, KH, KW, ph, pw, sh, sw, 1, 1, i1)
        port.unfold("col2im", col, N, C, H, W, Ho, Wo, KH, KW, ph, pw, sh, sw, 1, 1, i2)
        assert np.array_equal(i1, i2)


def test_host_geometry_matches_oracle(wf):
    import importlib
    cv = importlib.import_module("cl-waffe2_b200.conv")
    for (hin, k, pad, st) in [(28, 3, 0, 1), (12, 5, 0, 1), (16, 4, 0, 4), (9, 3, 1, 2), (7, 2, 1, 1), (11, 3, 0, 3)]:
        g = cv.Geometry((2, 3, hin, hin), (k, k), (st, st), (pad, pad))
        assert g.Ho == OC.conv_out_size(hin, pad, 1, k, st) == cv.conv_out_size(hin, pad, 1, k, st)
        assert cv.pool_out_size(hin, pad, k, st) == OC.pool_out_size(hin, pad, k, st)
        assert g.H == hin + 2 * pad + g.Ho % st
        s = g.struct(True)
        assert (s.N, s.C, s.H, s.W, s.Ho, s.Wo, s.KH, s.KW) == (2, 3, g.H, g.W, g.Ho, g.Wo, k, k)
        assert (g.struct(False).H, g.M, g.K) == (hin, 2 * g.Ho * g.Wo, 3 * k * k)
    assert cv._pair(3, "stride") == (3, 3) and cv._pair([1, 2], "stride") == (1, 2)
    with pytest.raises(ValueError):
        cv._pair((1, 2, 3), "stride")
