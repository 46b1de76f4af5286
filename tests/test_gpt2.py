"""GPT-2 forward (BASELINE.json config 5).  The reference has no tests or golden vectors for its
GPT-2 example (it needs downloaded weights), so the checker is oracle/gpt2.py — a numpy
restatement of examples/gpt-2/{model,utils}.lisp — in fp64; tolerance 1e-3 relative to max|h|
on the final hidden state (deep fp32 chain)."""
import importlib

import numpy as np
import pytest

from oracle import gpt2 as O

CFG = dict(n_vocab=211, n_ctx=48, n_emb=64, n_head=4, n_layer=3)


def _weights(cfg, seed=0):
    g = importlib.import_module("cl-waffe2_b200.gpt2")
    return g.init_weights(cfg, seed)


# ------------------------------------------------------------------ oracle self-checks (CPU)
def test_oracle_attention_is_normalised_over_heads():
    """(!softmax w :avoid-overflow nil) with the default axis=1 on a 4-D tensor: sums over HEADS are 1
    (examples/gpt-2/utils.lisp:236-242), not over keys."""
    rng = np.random.default_rng(0)
    w = rng.normal(size=(2, 4, 5, 5))
    s = O.softmax_no_shift(w, axis=1)
    np.testing.assert_allclose(s.sum(axis=1), 1.0, rtol=1e-12)
    assert not np.allclose(s.sum(axis=-1), 1.0)


def test_oracle_split_merge_heads_roundtrip():
    rng = np.random.default_rng(1)
    x = rng.normal(size=(2, 6, 3 * 16))
    q, k, v = (O.split_heads(x, w, 4) for w in "QKV")
    assert q.shape == (2, 4, 6, 4) and k.shape == (2, 4, 4, 6) and v.shape == (2, 4, 6, 4)
    np.testing.assert_array_equal(O.merge_heads(q), x[..., :16])
    np.testing.assert_array_equal(O.merge_heads(k.transpose(0, 1, 3, 2)), x[..., 16:32])


def test_oracle_layernorm_and_embedding():
    rng = np.random.default_rng(2)
    x = rng.normal(size=(3, 5, 32))
    y = O.layer_norm(x, np.ones(32), np.zeros(32))
    np.testing.assert_allclose(y.mean(-1), 0, atol=1e-12)
    np.testing.assert_allclose(y.var(-1), 1, rtol=1e-4)
    wte, wpe = rng.normal(size=(10, 4)), rng.normal(size=(7, 4))
    ctx = np.array([[0.0, 9.0, 3.2]])           # float ids, rounded (utils.lisp:60-75)
    e = O.embed(ctx, wte, wpe)
    np.testing.assert_array_equal(e[0, 2], wte[3] + wpe[2])


def test_oracle_fp32_tracks_fp64():
    W = _weights(CFG)
    ctx = np.random.default_rng(3).integers(0, CFG["n_vocab"], size=(2, 16)).astype(np.float32)
    h64 = O.forward(ctx, W, CFG, np.float64)
    h32 = O.forward(ctx, W, CFG, np.float32)
    assert np.abs(h32 - h64).max() / np.abs(h64).max() < 1e-4


# ------------------------------------------------------------------ device parity
@pytest.mark.gpu
@pytest.mark.parametrize("cols", [768, 64, 1024, 260, 12])
def test_layernorm_one_kernel_equals_node_chain(gpu, cols):
    """wf2cu_layernorm vs the ten-node LayerNorm-Revisit chain (examples/gpt-2/utils.lisp:114-127): the element-wise
    operations are the same in the same order, only the two row sums use another (shuffle) tree -> a few ulp; both
    within 1e-5 of fp64."""
    g = importlib.import_module("cl-waffe2_b200.gpt2")
    rng = np.random.default_rng(cols)
    X = (rng.normal(size=(3, 37, cols)) * 3 + 1.5).astype(np.float32)
    G = rng.normal(size=(cols,)).astype(np.float32)
    Bv = rng.normal(size=(cols,)).astype(np.float32)
    x, gg, bb = gpu.from_numpy(X), gpu.from_numpy(G), gpu.from_numpy(Bv)
    old = g.FUSE_LAYERNORM
    try:
        g.FUSE_LAYERNORM = True
        fused = g.layer_norm(x, gg, bb).to_numpy()
        g.FUSE_LAYERNORM = False
        chain = g.layer_norm(x, gg, bb).to_numpy()
    finally:
        g.FUSE_LAYERNORM = old
    truth = O.layer_norm(X.astype(np.float64), G.astype(np.float64), Bv.astype(np.float64))
    scale = np.abs(truth).max()
    assert np.abs(fused - truth).max() <= 1e-5 * scale
    assert np.abs(chain - truth).max() <= 1e-5 * scale
    np.testing.assert_allclose(fused, chain, rtol=0, atol=4e-6 * scale)


@pytest.mark.gpu
def test_layernorm_declines_unsupported_rows(gpu):
    lib = gpu._lib.load()
    x = gpu.make_tensor([4, 2050])
    v = gpu.make_tensor([2050])
    assert lib.wf2cu_layernorm(x.vec.ptr, 0, v.vec.ptr, 0, v.vec.ptr, 0, x.vec.ptr, 0, 4, 2050, 1e-5) == 3   # UNSUPPORTED


@pytest.fixture(params=[True, False], ids=["fused-layernorm", "node-chain"])
def ln_mode(request):
    g = importlib.import_module("cl-waffe2_b200.gpt2")
    old = g.FUSE_LAYERNORM
    g.FUSE_LAYERNORM = request.param
    yield request.param
    g.FUSE_LAYERNORM = old


@pytest.mark.gpu
@pytest.mark.parametrize("lm_head", [False, True])
def test_gpt2_forward_matches_oracle(gpu, lm_head, ln_mode):
    g = importlib.import_module("cl-waffe2_b200.gpt2")
    W = _weights(CFG, seed=5)
    ctx = np.random.default_rng(4).integers(0, CFG["n_vocab"], size=(3, 24)).astype(np.float32)
    model = g.GPT2(CFG, W)
    out = model(gpu.from_numpy(ctx), lm_head=lm_head).to_numpy()
    truth = O.forward(ctx, W, CFG, np.float64, lm_head=lm_head)
    assert out.shape == truth.shape
    err = np.abs(out - truth).max() / np.abs(truth).max()
    ref = np.abs(O.forward(ctx, W, CFG, np.float32, lm_head=lm_head) - truth).max() / np.abs(truth).max()
    assert err <= 1e-3, (err, ref)
    assert err <= max(50 * ref, 2e-5), (err, ref)


@pytest.mark.gpu
def test_gpt2_blocks_on_tensor_cores(gpu):
    """Same forward with the tcgen05 3xTF32 path forced for every !matmul."""
    g = importlib.import_module("cl-waffe2_b200.gpt2")
    cfg = dict(n_vocab=300, n_ctx=128, n_emb=128, n_head=4, n_layer=2)
    W = _weights(cfg, seed=6)
    ctx = np.random.default_rng(7).integers(0, cfg["n_vocab"], size=(2, 128)).astype(np.float32)
    lib = gpu._lib.load()
    gpu._lib.check(lib.wf2cu_gemm_set_mode(2))
    try:
        out = g.GPT2(cfg, W)(gpu.from_numpy(ctx)).to_numpy()
    finally:
        gpu._lib.check(lib.wf2cu_gemm_set_mode(0))
    truth = O.forward(ctx, W, cfg, np.float64)
    assert np.abs(out - truth).max() / np.abs(truth).max() <= 1e-3


# ------------------------------------------------------------------ the two attention kernels (csrc/attention.cu)
def _attention_truth(qkv, n_head):
    """examples/gpt-2/utils.lisp:222-244 in fp64: scores / N, exp, normalised over the HEADS axis, times V, merged heads."""
    x = qkv.astype(np.float64)
    Q, K, V = (O.split_heads(x, w, n_head) for w in "QKV")
    w = (Q @ K) / x.shape[1]
    p = O.softmax_no_shift(w, axis=1)
    return p, O.merge_heads(p @ V)


@pytest.mark.gpu
@pytest.mark.parametrize("B,N", [(1, 128), (2, 256), (1, 384), (3, 640)])
def test_attention_kernels_match_fp64_and_the_node_sequence(gpu, B, N):
    """wf2cu_attn_heads_softmax + wf2cu_attn_pv at 12 heads x 64 features (GPT-2 small) against fp64 and against the node-by-node
    sequence the example composes (the unfused path of gpt2.self_attention).  Bars: weights 1e-4 relative (one truncated TF32
    pass on the scores, divided by N before exp), output 1e-4 relative to max|out| (3xTF32 product)."""
    g = importlib.import_module("cl-waffe2_b200.gpt2")
    H, D = 12, 64
    rng = np.random.default_rng(B * 1000 + N)
    qkv = (rng.normal(size=(B, N, 3 * H * D)) * 2).astype(np.float32)
    x = gpu.from_numpy(qkv)
    lib = gpu._lib.load()
    p_dev = gpu.make_tensor([B, H, N, N])
    gpu._lib.check(lib.wf2cu_attn_heads_softmax(x.vec.ptr, 0, B, N, H, D, float(N), p_dev.vec.ptr, 0))
    o_dev = gpu.make_tensor([B, N, H * D])
    gpu._lib.check(lib.wf2cu_attn_pv(p_dev.vec.ptr, 0, x.vec.ptr, 0, B, N, H, D, o_dev.vec.ptr, 0))
    p, out = p_dev.to_numpy(), o_dev.to_numpy()
    p64, out64 = _attention_truth(qkv, H)
    np.testing.assert_allclose(p.sum(axis=1), 1.0, rtol=0, atol=1e-6)            # normalised over the heads axis
    rel_p = np.abs(p - p64).max() / np.abs(p64).max()
    rel_o = np.abs(out - out64).max() / np.abs(out64).max()
    assert rel_p <= 1e-4, rel_p
    assert rel_o <= 1e-4, rel_o
    # the product alone, on the device's own weights: 3xTF32 grade
    pv64 = O.merge_heads(p.astype(np.float64) @ O.split_heads(qkv.astype(np.float64), "V", H))
    assert np.abs(out - pv64).max() / np.abs(pv64).max() <= 2e-5
    # and the whole block against the node sequence
    old = g.FUSE_ATTENTION
    try:
        g.FUSE_ATTENTION = False
        nodes = g.self_attention(x, H).to_numpy()
        g.FUSE_ATTENTION = True
        fused = g.self_attention(x, H).to_numpy()
    finally:
        g.FUSE_ATTENTION = old
    np.testing.assert_array_equal(fused, out)
    assert np.abs(nodes - out64).max() / np.abs(out64).max() <= 1e-4
    assert np.abs(fused - nodes).max() / np.abs(out64).max() <= 1e-4


@pytest.mark.gpu
def test_attention_divisor_not_a_power_of_two_and_unsupported_shapes(gpu):
    """divisor = 384 takes the true-division path (`!div` is a divide, x * (1/384) would differ in the last bit); shapes the
    kernels do not take are declined with WF2CU_ERR_UNSUPPORTED so the host issues the nodes."""
    H, D, B, N = 12, 64, 1, 384
    rng = np.random.default_rng(9)
    qkv = (rng.normal(size=(B, N, 3 * H * D)) * 2).astype(np.float32)
    x = gpu.from_numpy(qkv)
    lib = gpu._lib.load()
    p_dev = gpu.make_tensor([B, H, N, N])
    gpu._lib.check(lib.wf2cu_attn_heads_softmax(x.vec.ptr, 0, B, N, H, D, 384.0, p_dev.vec.ptr, 0))
    p64, _ = _attention_truth(qkv, H)
    assert np.abs(p_dev.to_numpy() - p64).max() / np.abs(p64).max() <= 1e-4
    y = gpu.make_tensor([1, 100, 3 * 64])
    q = gpu.make_tensor([1, 4, 100, 100])
    assert lib.wf2cu_attn_heads_softmax(y.vec.ptr, 0, 1, 100, 4, 16, 100.0, q.vec.ptr, 0) == 3
    assert lib.wf2cu_attn_pv(q.vec.ptr, 0, y.vec.ptr, 0, 1, 100, 4, 16, y.vec.ptr, 0) == 3
    assert lib.wf2cu_attn_heads_softmax(x.vec.ptr, 0, B, 200, H, D, 200.0, p_dev.vec.ptr, 0) == 3     # seq not a multiple of 128


@pytest.mark.gpu
@pytest.mark.parametrize("B,N", [(1, 128), (2, 384), (3, 1024), (19, 512)])     # (19, 512): 152 six-head work items, so some CTAs take two
def test_one_call_attention_equals_the_two_launches(gpu, B, N):
    """wf2cu_attn_fused (one pass over the heads: un-normalised exponentials + reciprocal denominators in the library's workspace,
    p = RN(e r) formed by the P.V kernel) against wf2cu_attn_heads_softmax + wf2cu_attn_pv: the same products in the same order,
    so the merged-heads result is bit-identical; and both against the oracle's SelfAttention."""
    import ctypes as C
    lib = gpu._lib.load()
    H, D = 12, 64
    rng = np.random.default_rng(100 + N)
    qkv = rng.normal(0, 1.0, size=(B, N, 3 * H * D)).astype(np.float32)
    g = gpu.from_numpy(qkv)
    P = gpu.make_tensor((B, H, N, N))
    two = gpu.make_tensor((B, N, H * D))
    one = gpu.make_tensor((B, N, H * D))
    gpu._lib.check(lib.wf2cu_attn_heads_softmax(g.vec.ptr, 0, B, N, H, D, float(N), P.vec.ptr, 0))
    gpu._lib.check(lib.wf2cu_attn_pv(P.vec.ptr, 0, g.vec.ptr, 0, B, N, H, D, two.vec.ptr, 0))
    for _ in range(2):                                   # twice: the workspace is reused, nothing may depend on its old contents
        gpu._lib.check(lib.wf2cu_attn_fused(g.vec.ptr, 0, B, N, H, D, float(N), one.vec.ptr, 0))
    a, b = one.to_numpy(), two.to_numpy()
    np.testing.assert_array_equal(a, b)
    # fp64 restatement of examples/gpt-2/utils.lisp:222-244 (scores / N, softmax over the HEADS axis, no mask)
    x = qkv.astype(np.float64).reshape(B, N, 3, H, D)
    q, k, v = (np.transpose(x[:, :, i], (0, 2, 1, 3)) for i in range(3))
    e = np.exp(q @ np.swapaxes(k, -1, -2) / N)
    want = np.transpose((e / e.sum(axis=1, keepdims=True)) @ v, (0, 2, 1, 3)).reshape(B, N, H * D)
    assert np.abs(a - want).max() <= 2e-5 * np.abs(want).max()


@pytest.mark.gpu
def test_one_call_attention_repeated_behind_other_work(gpu):
    """wf2cu_attn_fused overlaps its V split with the score kernel through programmatic dependent launch (the split does not wait at its
    start; one of its CTAs waits for the score kernel before it exits, which is what orders P.V behind BOTH).  A missing edge there would
    show as a P.V kernel that reads E or the V planes too early — only sometimes.  30 calls, each queued behind a different amount of
    unrelated work and each overwriting the result buffer first, must all equal the two-launch result bit for bit."""
    lib = gpu._lib.load()
    B, N, H, D = 4, 512, 12, 64
    rng = np.random.default_rng(7)
    qkv = gpu.from_numpy(rng.normal(0, 1.0, size=(B, N, 3 * H * D)).astype(np.float32))
    P = gpu.make_tensor((B, H, N, N))
    two = gpu.make_tensor((B, N, H * D))
    one = gpu.make_tensor((B, N, H * D))
    big = gpu.from_numpy(rng.normal(size=(2048, 2048)).astype(np.float32))
    gpu._lib.check(lib.wf2cu_attn_heads_softmax(qkv.vec.ptr, 0, B, N, H, D, float(N), P.vec.ptr, 0))
    gpu._lib.check(lib.wf2cu_attn_pv(P.vec.ptr, 0, qkv.vec.ptr, 0, B, N, H, D, two.vec.ptr, 0))
    want = two.to_numpy()
    for i in range(30):
        gpu.ops.fill_(one, float(i))
        for _ in range(i % 4):
            gpu.ops.exp(big)
        if i % 5 == 4:
            gpu.ops.matmul(big, big)
        gpu._lib.check(lib.wf2cu_attn_fused(qkv.vec.ptr, 0, B, N, H, D, float(N), one.vec.ptr, 0))
        np.testing.assert_array_equal(one.to_numpy(), want, err_msg=f"call {i}")
