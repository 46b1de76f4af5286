"""MLP training-step benchmark (BASELINE.json config 4: MNIST MLP 784-256-256-10, softmax-cross-entropy,
Adam, batch 1024 synthetic inputs, batch-sharded over 1/2/4/8 B200 with a NCCL gradient allreduce).
Called from bench.py; returns a dict for the `mlp` key of the JSON line."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib, nn

DIMS = [784, 256, 256, 10]


def _comm_init(dist, rank: int, world: int):
    """One NCCL communicator per process (one process per GPU); the 128-byte id travels over the
    torch.distributed plumbing."""
    import torch
    lib = _lib.load()
    buf = (C.c_ubyte * 128)()
    if rank == 0:
        _lib.check(lib.wf2cu_comm_unique_id(buf))
    t = torch.tensor(list(buf), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, src=0)
    ids = bytes(t.cpu().tolist())
    _lib.check(lib.wf2cu_comm_init(rank, world, C.create_string_buffer(ids, 128)))


def _stage(msg):
    import os, sys, time
    if os.environ.get("WF2_BENCH_TRACE"):
        print(f"[mlp rank {os.environ.get('RANK', '0')} t={time.time() % 1000:.1f}] {msg}", file=sys.stderr, flush=True)


def _all_gather_bytes(dist, world):
    def f(b: bytes):
        import torch
        t = torch.tensor(list(b), dtype=torch.uint8, device="cuda")
        outs = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(outs, t)
        return [bytes(o.cpu().tolist()) for o in outs]
    return f


def _run_one(dist, rank, world, per_gpu_batch, steps, warmup, seed=0, exchange="nccl", workload="mlp"):
    lib = _lib.load()
    _stage(f"run_one {workload} batch={per_gpu_batch}")
    chk = _lib.check
    rng = np.random.default_rng(seed + 17 * rank)          # each rank owns a different shard
    if workload == "cnn":
        from . import conv
        X = rng.uniform(0, 1, size=(per_gpu_batch, 1, 28, 28)).astype(np.float32)
        model = conv.MNISTCNN(seed=1234)
    else:
        X = rng.uniform(0, 1, size=(per_gpu_batch, DIMS[0])).astype(np.float32)
        model = nn.MLPSequence(DIMS, "relu", seed=1234)    # same seed everywhere: replicas start identical
    Y = np.eye(DIMS[-1], dtype=np.float32)[rng.integers(DIMS[-1], size=per_gpu_batch)]
    step = nn.CompiledStep(model, "softmax-cross-entropy-sum", nn.Adam(model.arena, lr=1e-3), X.shape, Y.shape, world=world,
                           exchange=exchange)
    if world > 1 and exchange == "p2p":
        nn.p2p_attach(model.arena, rank, world, _all_gather_bytes(dist, world))
    step.set_inputs(X, Y)

    def barrier():
        chk(lib.wf2cu_sync())
        if dist is not None:
            dist.barrier()

    def max_ranks(v):
        if dist is None:
            return v
        import torch
        t = torch.tensor([v], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for i in range(max(warmup, 3)):
        step.step()
        if i < 3:
            chk(lib.wf2cu_sync())
            _stage(f"warm-up step {i} done")
    first_loss = None
    barrier()
    l0 = C.c_int64(); l1 = C.c_int64()
    chk(lib.wf2cu_launch_count(C.byref(l0)))
    chk(lib.wf2cu_timer_start())
    for _ in range(steps):
        step.step()
    ms = C.c_float()
    chk(lib.wf2cu_timer_stop(C.byref(ms)))
    chk(lib.wf2cu_launch_count(C.byref(l1)))
    barrier()
    dev_ms = max_ranks(ms.value) / steps
    loss_after = step.loss_value()

    # e2e: the batch comes from pinned host memory every step and the loss goes back to the host
    hx = C.c_void_p(); hy = C.c_void_p()
    chk(lib.wf2cu_host_alloc(X.nbytes, C.byref(hx))); chk(lib.wf2cu_host_alloc(Y.nbytes, C.byref(hy)))
    C.memmove(hx.value, X.ctypes.data, X.nbytes); C.memmove(hy.value, Y.ctypes.data, Y.nbytes)
    out = C.c_float()

    # Two input buffers: the batch of step i+1 is copied on a side stream (wf2cu_branch_*) while step i trains.
    # The captured graph reads step.x / step.y, so the side stream fills a staging pair and a device-to-device
    # copy (3 MB at HBM speed) moves it in front of the step.  Exactly `steps` uploads inside the timed region.
    from .tensor import make_tensor
    sx, sy = make_tensor(list(X.shape)), make_tensor(list(Y.shape))

    def upload():
        chk(lib.wf2cu_h2d(sx.vec.ptr, 0, hx, X.nbytes))
        chk(lib.wf2cu_h2d(sy.vec.ptr, 0, hy, Y.nbytes))

    def e2e_run(k):
        upload()
        for i in range(k):
            chk(lib.wf2cu_d2d(step.x.vec.ptr, 0, sx.vec.ptr, 0, X.nbytes))
            chk(lib.wf2cu_d2d(step.y.vec.ptr, 0, sy.vec.ptr, 0, Y.nbytes))
            if i + 1 < k:
                chk(lib.wf2cu_branch_begin(1))
                try:
                    upload()
                finally:
                    chk(lib.wf2cu_branch_end(1))
            step.step()
            chk(lib.wf2cu_get_f32(step.loss.vec.ptr, step.loss.lower().offset, C.byref(out)))
            if i + 1 < k:
                chk(lib.wf2cu_branch_join(1))

    e2e_run(3)
    barrier()
    chk(lib.wf2cu_timer_start())
    e2e_run(steps)
    chk(lib.wf2cu_timer_stop(C.byref(ms)))
    barrier()
    e2e_ms = max_ranks(ms.value) / steps
    chk(lib.wf2cu_host_free(hx)); chk(lib.wf2cu_host_free(hy))
    gb = per_gpu_batch * world
    step.close()
    p2p_timeout = False
    if world > 1 and exchange == "p2p":
        p2p_timeout = nn.p2p_timed_out()
        barrier()
        nn.p2p_detach()
    # replicas must stay identical: checksum of the parameter arena, compared across ranks
    psum = float(np.abs(model.arena.params[0].to_numpy()).sum())
    if dist is not None:
        import torch
        t = torch.tensor([psum, -psum], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        in_sync = bool(t[0].item() == -t[1].item())
    else:
        in_sync = True
    return {"exchange": exchange if world > 1 else None, "replicas_identical": in_sync, "p2p_timed_out": p2p_timeout,"per_gpu_batch": per_gpu_batch, "global_batch": gb, "ms_per_step": dev_ms,
            "samples_per_s": gb / (dev_ms * 1e-3), "e2e_ms_per_step": e2e_ms, "e2e_samples_per_s": gb / (e2e_ms * 1e-3),
            "h2d_bytes_per_step": int(X.nbytes + Y.nbytes), "d2h_bytes_per_step": 4,
            "launches_per_step": (l1.value - l0.value) / steps, "loss_after": loss_after,
            "pool_bytes": step.pool.nbytes()}


def run(wf, dist, rank: int, world: int, steps: int = 200, warmup: int = 20, exchange: str = "p2p") -> dict:
    if world > 1:
        _stage("comm init")
        _comm_init(dist, rank, world)
        _stage("comm ready")
    res = {"model": "MNIST MLP 784-256-256-10, relu, softmax-cross-entropy (!sum), Adam lr 1e-3, fp32",
           "data": "synthetic U(0,1) inputs, one-hot labels", "n_gpus": world,
           "collective": ("one kernel per rank inside the CUDA graph: sum of all ranks' 269,322-float gradient arenas over "
                          "NVLink peer memory (CUDA IPC) + 1/world + Adam (csrc/p2p.cu); weak_nccl = the same step with "
                          "ncclAllReduce + Adam" if exchange == "p2p" else
                          "one NCCL allreduce(sum) of the 269,322-float gradient bucket per step, inside the CUDA graph")
           if world > 1 else None}
    res["weak"] = _run_one(dist, rank, world, 1024, steps, warmup, exchange=exchange)                 # 1024 per GPU
    if world > 1 and 1024 % world == 0:
        res["strong"] = _run_one(dist, rank, world, 1024 // world, steps, warmup, exchange=exchange)  # 1024 global
    if world > 1 and exchange == "p2p":
        res["weak_nccl"] = _run_one(dist, rank, world, 1024, steps, warmup, exchange="nccl")          # the library baseline
    if world > 1:
        _lib.check(_lib.load().wf2cu_comm_destroy())
    return res


def run_cnn(wf, steps: int = 100, warmup: int = 10) -> dict:
    """§8 f-3: the MNIST-CNN training step (examples/mnist/cnn.lisp:5-26; Conv2D 1->4 3x3, relu, MaxPool 2,
    Conv2D 4->16 5x5, relu, MaxPool 2, LinearLayer 256->10, softmax-cross-entropy (!sum), Adam) on one GPU, at
    the example's batch of 100 and at 1024."""
    res = {"model": "MNIST-CNN (examples/mnist/cnn.lisp): Conv2D(1,4,3x3) relu MaxPool2D(2) Conv2D(4,16,5x5) relu "
                    "MaxPool2D(2) LinearLayer(256,10), softmax-cross-entropy (!sum), Adam lr 1e-3, fp32",
           "data": "synthetic U(0,1) images (N 1 28 28), one-hot labels"}
    for b in (100, 1024):
        res[f"batch{b}"] = _run_one(None, 0, 1, b, steps, warmup, workload="cnn")
    return res
