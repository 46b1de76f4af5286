"""GPU parity of the MLP training step (configs C1 and C4): the compiled CUDA-graph step
against the oracle's node-by-node restatement on the same host-generated weights and data.
Tolerance: per-step loss within 1e-4 relative (the GEMM bar) over the compared steps."""
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _nn(gpu):
    return importlib.import_module("cl-waffe2_b200.nn")


def test_c1_sin_wave_sgd_matches_oracle(gpu):
    # examples/mlp_sin_wave.lisp:41-78 — 100 -> 10 -> 1, sigmoid, MSE(mean), SGD lr 1e-3, batch 100
    nn = _nn(gpu)
    from oracle.mlp import OracleMLP
    i = np.arange(100 * 100, dtype=np.float32)
    X = np.sin(np.float32(0.01) * i + np.float32(0.1)).reshape(100, 100).astype(np.float32)
    Y = np.cos(np.float32(0.01) * np.arange(100, dtype=np.float32) + np.float32(0.1)).reshape(100, 1).astype(np.float32)
    init = nn.mlp_init_arrays([100, 10, 1], seed=1)
    model = nn.MLPSequence([100, 10, 1], "sigmoid", init=init)
    step = nn.CompiledStep(model, "mse-mean", nn.SGD(model.arena, lr=1e-3), X.shape, Y.shape)
    step.set_inputs(X, Y)
    ref = OracleMLP([100, 10, 1], "sigmoid", init, "mse-mean", opt="sgd", lr=1e-3)
    ours, theirs = [], []
    for _ in range(30):
        step.step()
        ours.append(step.loss_value())
        theirs.append(ref.step(X, Y))
    np.testing.assert_allclose(ours, theirs, rtol=1e-4)
    assert ours[-1] < ours[0]                     # nn/t/regression.lisp:429-454: loss decreases
    assert step.graph is not None                 # replayed as a CUDA graph from the 2nd step on
    for p, q in zip(model.parameters(), [ref.W[0], ref.b[0], ref.W[1], ref.b[1]]):
        np.testing.assert_allclose(p.to_numpy().ravel(), q.vec, rtol=1e-4, atol=1e-6)


@pytest.fixture(params=[True, False], ids=["fused", "node-by-node"])
def fuse(request, gpu):
    """Both instruction sequences: with the backend's node fusions (GEMM epilogues, one-kernel loss) and
    one launch per node."""
    nn = _nn(gpu)
    old = (nn.FUSE, nn.CONCURRENT)
    nn.FUSE = nn.CONCURRENT = request.param      # fused + side-stream branches  vs  strictly in-order, one launch per node
    yield request.param
    nn.FUSE, nn.CONCURRENT = old


@pytest.mark.parametrize("use_graph", [True, False])
def test_c4_mnist_mlp_adam_matches_oracle(gpu, use_graph, fuse):
    # examples/mnist/mlp.lisp:12-36 — 784-256-256-10, relu, softmax-cross-entropy (!sum), Adam lr 1e-3
    nn = _nn(gpu)
    from oracle.mlp import OracleMLP
    rng = np.random.default_rng(0)
    B = 64
    X = rng.uniform(0, 1, size=(B, 784)).astype(np.float32)
    Y = np.eye(10, dtype=np.float32)[rng.integers(10, size=B)]
    dims = [784, 256, 256, 10]
    init = nn.mlp_init_arrays(dims, seed=2)
    model = nn.MLPSequence(dims, "relu", init=init)
    step = nn.CompiledStep(model, "softmax-cross-entropy-sum", nn.Adam(model.arena, lr=1e-3), X.shape, Y.shape,
                           use_graph=use_graph)
    step.set_inputs(X, Y)
    ref = OracleMLP(dims, "relu", init, "softmax-cross-entropy-sum", opt="adam", lr=1e-3)
    ours, theirs = [], []
    for _ in range(12):
        step.step()
        ours.append(step.loss_value())
        theirs.append(ref.step(X, Y))
    np.testing.assert_allclose(ours, theirs, rtol=1e-4)
    assert ours[-1] < 0.7 * ours[0]


def test_gradients_match_oracle(gpu, fuse):
    nn = _nn(gpu)
    from oracle.mlp import OracleMLP
    rng = np.random.default_rng(3)
    dims = [40, 24, 8]
    X = rng.normal(size=(16, 40)).astype(np.float32)
    Y = np.eye(8, dtype=np.float32)[rng.integers(8, size=16)]
    init = nn.mlp_init_arrays(dims, seed=4)
    model = nn.MLPSequence(dims, "relu", init=init)
    opt = nn.SGD(model.arena, lr=0.0)             # lr 0: parameters stay put, gradients are what we compare
    step = nn.CompiledStep(model, "softmax-cross-entropy-sum", opt, X.shape, Y.shape, use_graph=False)
    step.set_inputs(X, Y)
    step.step()
    ref = OracleMLP(dims, "relu", init, "softmax-cross-entropy-sum", opt="sgd", lr=0.0)
    loss, gW, gb = ref.forward_backward(X, Y)
    assert abs(step.loss_value() - loss) <= 1e-4 * abs(loss)
    got = [p.grad.to_numpy().ravel() for p in model.parameters()]
    exp = [gW[0].vec, gb[0].vec, gW[1].vec, gb[1].vec]
    for g, e in zip(got, exp):
        np.testing.assert_allclose(g, e, rtol=1e-4, atol=1e-6)


def test_gemm_epilogues(gpu):
    """wf2cu_gemm_ex: + bias, !relu, and the relu-backward mask, on the SIMT split-K path and (forced) on the
    tensor path with its separate epilogue pass; against numpy with the reference's relu = x * (x > 0)."""
    import ctypes as C
    wf = gpu
    rng = np.random.default_rng(5)
    lib = wf._lib.load()
    for (m, n, k), mode in [((100, 70, 33), 0), ((1024, 256, 784), 0), ((256, 512, 128), 2)]:
        A = rng.normal(size=(m, k)).astype(np.float32)
        W = rng.normal(size=(n, k)).astype(np.float32)
        b = rng.normal(size=(n,)).astype(np.float32)
        M = rng.normal(size=(m, n)).astype(np.float32)
        ga, gw, gb, gm = (wf.from_numpy(a) for a in (A, W, b, M))
        wf._lib.check(lib.wf2cu_gemm_set_mode(mode))
        try:
            pre = (A.astype(np.float64) @ W.T.astype(np.float64))
            tol = dict(rtol=2e-4, atol=2e-4 * np.abs(pre).max())
            r = wf.ops.matmul(ga, wf.ops.t(gw), bias=gb).to_numpy()
            np.testing.assert_allclose(r, pre + b, **tol)
            r = wf.ops.matmul(ga, wf.ops.t(gw), bias=gb, relu=True).to_numpy()
            np.testing.assert_allclose(r, np.maximum(pre + b, 0), **tol)
            r = wf.ops.matmul(ga, wf.ops.t(gw), relu_mask=gm).to_numpy()
            np.testing.assert_allclose(r, pre * (M > 0), **tol)
        finally:
            wf._lib.check(lib.wf2cu_gemm_set_mode(0))


@pytest.mark.gpu
def test_p2p_fused_allreduce_adam_two_ranks():
    """csrc/p2p.cu on 2 GPUs (skipped on a 1-GPU box): the fused peer-memory reduce-scatter + Adam + all-gather
    kernel is bitwise equal to a numpy emulation, equal to the NCCL route at world 2, replicas bit-identical."""
    import subprocess, sys, os
    import ctypes as C
    wf = importlib.import_module("cl-waffe2_b200")
    n = C.c_int(0)
    wf._lib.check(wf._lib.load().wf2cu_device_count(C.byref(n)))
    if n.value < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(root, "scripts", "p2p_check.py")],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
