"""World-size-2 test of the batch-sharded training step's HOST logic on CPU (gloo): shard
partition, flat gradient bucket layout, allreduce(sum) x 1/world, identical optimizer on every rank.
The per-rank forward/backward is the oracle MLP (oracle/mlp.py); on the GPU box the same protocol
runs with wf2cu_allreduce_sum_f32 (NCCL) inside the step's CUDA graph (cl-waffe2_b200/nn.py
CompiledStep, world > 1)."""
import importlib
import os
import socket

import numpy as np
import pytest

DIMS = [20, 16, 16, 5]
GLOBAL_BATCH = 32
STEPS = 4


def _data():
    rng = np.random.default_rng(0)
    X = rng.uniform(0, 1, size=(GLOBAL_BATCH, DIMS[0])).astype(np.float32)
    Y = np.eye(DIMS[-1], dtype=np.float32)[rng.integers(DIMS[-1], size=GLOBAL_BATCH)]
    return X, Y


def _flatten(nn, grads, shapes):
    offs, total = nn.bucket_layout(shapes)
    b = np.zeros(total, np.float32)
    for g, o in zip(grads, offs):
        b[o:o + g.size] = g.ravel()
    return b


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nn = importlib.import_module("cl-waffe2_b200.nn")
    from oracle import mlp as OM
    from oracle import loopnest as L
    X, Y = _data()
    lo, hi = nn.shard_rows(GLOBAL_BATCH, rank, world)
    m = OM.OracleMLP(DIMS, "relu", nn.mlp_init_arrays(DIMS, seed=7), "softmax-cross-entropy-sum")
    shapes = [p.shape for p in m.params()]
    offs, total = nn.bucket_layout(shapes)
    for _ in range(STEPS):
        _, gW, gb = m.forward_backward(X[lo:hi], Y[lo:hi])
        bucket = torch.from_numpy(_flatten(nn, [g.vec for g in gW] + [g.vec for g in gb], shapes))
        dist.all_reduce(bucket, op=dist.ReduceOp.SUM)            # the one collective of the step
        red = bucket.numpy()
        for g, o in zip(gW + gb, offs):
            g.vec[:] = red[o:o + g.vec.size]
        m.apply(gW, gb, grad_scale=1.0 / world)                  # x 1/world folded into the optimizer
    q.put((rank, np.concatenate([p.vec.ravel() for p in m.params()])))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_bucket_layout_and_shards():
    nn = importlib.import_module("cl-waffe2_b200.nn")
    shapes = [(256, 784), (256,), (256, 256), (256,), (10, 256), (10,)]      # config 4
    offs, total = nn.bucket_layout(shapes)
    assert offs == [0, 200704, 200960, 266496, 266752, 269312]
    assert total == 269324                        # 269 322 floats + 2 of alignment padding after b3
    assert all(o % 4 == 0 for o in offs)
    assert [nn.shard_rows(1024, r, 8) for r in (0, 7)] == [(0, 128), (896, 1024)]
    with pytest.raises(ValueError):
        nn.shard_rows(1000, 0, 3)


def test_two_rank_gloo_step_matches_single_process():
    import torch.multiprocessing as mp
    nn = importlib.import_module("cl-waffe2_b200.nn")
    from oracle import mlp as OM
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # replicas stay bit-identical (no broadcast after step 0)
    np.testing.assert_array_equal(res[0], res[1])
    # and equal the single-process step on the global batch: sce-diff-static divides by the LOCAL
    # batch (nn/criterion.lisp:100-106), so mean over ranks == the global-batch gradient
    X, Y = _data()
    m = OM.OracleMLP(DIMS, "relu", nn.mlp_init_arrays(DIMS, seed=7), "softmax-cross-entropy-sum")
    for _ in range(STEPS):
        m.step(X, Y)
    single = np.concatenate([p.vec.ravel() for p in m.params()])
    np.testing.assert_allclose(res[0], single, rtol=2e-4, atol=2e-6)
