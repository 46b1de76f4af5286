"""MLP training-step benchmark (BASELINE.json config 4: MNIST MLP 784-256-256-10, softmax-cross-entropy,
Adam, batch 1024 synthetic inputs, batch-sharded over 1/2/4/8 B200 with a gradient exchange over NVLink).
Called from bench.py; `run` returns (compact, detail): `compact` is the `mlp` object that ends the JSON line
(<= 1200 characters, so it survives the driver's 1500-character tail at every N), `detail` goes to `mlp_detail`.

At world > 1 the run also carries its own proof (`p2p_check`), because the driver's GPU test box has one GPU:
  * kernel_ulp_vs_emulation  the fused exchange+Adam kernel on synthetic arenas against a numpy fp32 emulation
                             (rank-ordered sum, x 1/world, Adam in the kernel's operation order): 0 = bitwise equal;
  * replicas_identical       parameter arenas bit-identical across ranks after every timed run;
  * loss_first5_vs_oracle    the first 5 losses of a sharded run (128 rows per rank) against the CPU checker's global step
                             (per-shard gradients summed x 1/world, `nn/criterion.lisp:100-106`), max relative difference;
  * loss_first20_p2p_vs_nccl max relative difference of the first 20 losses between the two exchanges — they differ
                             only in summation order (p2p: rank order; NCCL: its ring / tree order), so this is the
                             number behind any later gap between the two `loss_after` values."""
from __future__ import annotations

import ctypes as C
import time

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
    import os, sys
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


def shard_data(rank: int, per_gpu_batch: int, seed: int = 0, workload: str = "mlp"):
    """The synthetic shard rank `rank` owns (any rank can regenerate any shard: the oracle check needs them all)."""
    rng = np.random.default_rng(seed + 17 * rank)
    if workload == "cnn":
        X = rng.uniform(0, 1, size=(per_gpu_batch, 1, 28, 28)).astype(np.float32)
    else:
        X = rng.uniform(0, 1, size=(per_gpu_batch, DIMS[0])).astype(np.float32)
    Y = np.eye(DIMS[-1], dtype=np.float32)[rng.integers(DIMS[-1], size=per_gpu_batch)]
    return X, Y


def _run_one(dist, rank, world, per_gpu_batch, steps, warmup, seed=0, exchange="nccl", workload="mlp", e2e=True):
    lib = _lib.load()
    _stage(f"run_one {workload} batch={per_gpu_batch} world={world} exchange={exchange}")
    chk = _lib.check
    X, Y = shard_data(rank, per_gpu_batch, seed, workload)     # each rank owns a different shard
    if workload == "cnn":
        from . import conv
        model = conv.MNISTCNN(seed=1234)
    else:
        model = nn.MLPSequence(DIMS, "relu", seed=1234)        # same seed everywhere: replicas start identical
    step = nn.CompiledStep(model, "softmax-cross-entropy-sum", nn.Adam(model.arena, lr=1e-3), X.shape, Y.shape, world=world,
                           exchange=exchange)
    if world > 1 and exchange == "p2p":
        nn.p2p_attach(model.arena, rank, world, _all_gather_bytes(dist, world))
        if dist is not None:
            dist.barrier()                                      # nobody steps before every rank has mapped its peers
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

    first_losses = []
    for i in range(max(warmup, 3)):
        step.step()
        if i < 20:
            first_losses.append(step.loss_value())              # untimed: one 4-byte D2H + sync per warm-up step
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

    e2e_ms = None
    if e2e:
        # e2e: the batch comes from pinned host memory every step and the loss goes back to the host.
        # Two input buffers: the batch of step i+1 is copied on a side stream (wf2cu_branch_*) while step i trains.
        # The captured graph reads step.x / step.y, so the side stream fills a staging pair and a device-to-device
        # copy (3 MB at HBM speed) moves it in front of the step.  Exactly `steps` uploads inside the timed region.
        hx = C.c_void_p(); hy = C.c_void_p()
        chk(lib.wf2cu_host_alloc(X.nbytes, C.byref(hx))); chk(lib.wf2cu_host_alloc(Y.nbytes, C.byref(hy)))
        C.memmove(hx.value, X.ctypes.data, X.nbytes); C.memmove(hy.value, Y.ctypes.data, Y.nbytes)
        out = C.c_float()
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
    # replicas must stay identical: every rank's parameter arena, bit for bit, against rank 0's
    flat = np.concatenate([p.to_numpy().ravel() for p in model.arena.params])
    if dist is not None:
        import torch
        t = torch.from_numpy(flat).cuda()
        ref = t.clone()
        dist.broadcast(ref, src=0)
        same = torch.tensor([1 if torch.equal(t, ref) else 0], device="cuda")
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        in_sync = bool(same.item())
    else:
        in_sync = True
    return {"exchange": exchange if world > 1 else None, "replicas_identical": in_sync, "p2p_timed_out": p2p_timeout,
            "per_gpu_batch": per_gpu_batch, "global_batch": gb, "ms_per_step": dev_ms,
            "samples_per_s": gb / (dev_ms * 1e-3), "e2e_ms_per_step": e2e_ms,
            "e2e_samples_per_s": gb / (e2e_ms * 1e-3) if e2e_ms else None,
            "h2d_bytes_per_step": int(X.nbytes + Y.nbytes), "d2h_bytes_per_step": 4,
            "launches_per_step": (l1.value - l0.value) / steps, "loss_after": loss_after, "loss_first20": first_losses,
            "pool_bytes": step.pool.nbytes()}


# ------------------------------------------------------------------------------- the multi-rank proof
def _p2p_kernel_check(dist, rank: int, world: int) -> int:
    """The fused kernel alone on synthetic arenas every rank can regenerate: the result must equal a numpy fp32
    emulation (rank-ordered sum, x 1/world, Adam formula in the kernel's operation order) BITWISE.  Returns the
    largest difference in units in the last place over 5 steps on a ragged arena (0 = bitwise equal)."""
    f = np.float32
    n = 100003                                   # ragged: not a multiple of 4 or 256
    arena = nn.ParameterArena([(n,)])
    opt = nn.Adam(arena, lr=1e-3)
    p0 = np.random.default_rng(7).normal(size=n).astype(f)
    arena.params[0].copy_from_numpy(p0)
    nn.p2p_attach(arena, rank, world, _all_gather_bytes(dist, world))
    dist.barrier()
    lib = _lib.load()
    p, m, v = p0.copy(), np.zeros(n, f), np.zeros(n, f)
    worst = 0
    for it in range(1, 6):
        gs = [np.random.default_rng(1000 * it + r).normal(size=n).astype(f) * f(10.0 ** (-(r % 4))) for r in range(world)]
        arena.params[0].grad.copy_from_numpy(gs[rank])
        _lib.check(lib.wf2cu_sync())
        dist.barrier()
        opt.pre_step()
        opt.step_p2p()
        _lib.check(lib.wf2cu_sync())
        g = gs[0].copy()
        for r in range(1, world):
            g = (g + gs[r]).astype(f)
        g = (g * f(1.0 / world)).astype(f)
        lr_t = f(opt.lr_at(it)); omb1 = f(1.0) - f(opt.b1); omb2 = f(1.0) - f(opt.b2); eps = f(opt.eps)
        m = (m + ((g - m) * omb1)).astype(f)
        v = (v + (omb2 * ((g * g) - v))).astype(f)
        p = (p - ((m * lr_t) / (eps + np.sqrt(v)))).astype(f)
        got = arena.params[0].to_numpy()
        worst = max(worst, int(np.abs(got.view(np.int32).astype(np.int64) - p.view(np.int32).astype(np.int64)).max()))
        dist.barrier()
    if nn.p2p_timed_out():
        worst = -1
    dist.barrier()
    nn.p2p_detach()
    import torch
    t = torch.tensor([worst if worst >= 0 else 1 << 30], device="cuda", dtype=torch.int64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return int(t.item())


def _oracle_check(dist, rank: int, world: int, exchange: str, host_losses, per_gpu_batch: int = 128, steps: int = 5) -> float:
    """First `steps` global losses of a sharded run against the host checker's global step (`host_losses(world, per_gpu_batch,
    steps)`, supplied by bench.py, which alone may use the CPU checker); max relative difference."""
    import torch
    lib = _lib.load()
    X, Y = shard_data(rank, per_gpu_batch, 0)
    model = nn.MLPSequence(DIMS, "relu", seed=1234)
    step = nn.CompiledStep(model, "softmax-cross-entropy-sum", nn.Adam(model.arena, lr=1e-3), X.shape, Y.shape, world=world,
                           exchange=exchange)
    if exchange == "p2p":
        nn.p2p_attach(model.arena, rank, world, _all_gather_bytes(dist, world))
        dist.barrier()
    step.set_inputs(X, Y)
    mine = []
    for _ in range(steps):
        step.step()
        mine.append(step.loss_value())
    step.close()
    _lib.check(lib.wf2cu_sync())
    dist.barrier()
    if exchange == "p2p":
        nn.p2p_detach()
    t = torch.tensor(mine, device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    got = t.cpu().numpy()
    rel = -1.0
    if rank == 0:
        want = np.array(host_losses(world, per_gpu_batch, steps))
        rel = float(np.max(np.abs(got - want) / np.abs(want)))
    r = torch.tensor([rel], device="cuda", dtype=torch.float64)
    dist.broadcast(r, src=0)
    return float(r.item())


def _r(x, nd=1):
    return None if x is None else round(float(x), nd)


def run(wf, dist, rank: int, world: int, steps: int = 200, warmup: int = 20, exchange: str = "p2p", host_losses=None, cpu_fn=None):
    """Returns (compact, detail).  host_losses / cpu_fn: the CPU checker's global-step losses and the timed CPU baseline,
    both supplied by bench.py (nothing in this package touches the checker)."""
    lib = _lib.load()
    if world > 1:
        _stage("comm init")
        _comm_init(dist, rank, world)
        _stage("comm ready")
    detail = {"model": "MNIST MLP 784-256-256-10, relu, softmax-cross-entropy (!sum), Adam lr 1e-3, fp32",
              "data": "synthetic U(0,1) inputs, one-hot labels", "n_gpus": world,
              "collective": ("one kernel per rank inside the CUDA graph: sum of all ranks' 269,322-float gradient arenas over "
                             "NVLink peer memory (CUDA IPC) + 1/world + Adam (csrc/p2p.cu); weak_nccl = the same step with "
                             "ncclAllReduce + Adam" if exchange == "p2p" else
                             "one NCCL allreduce(sum) of the 269,322-float gradient bucket per step, inside the CUDA graph")
              if world > 1 else None}
    # a 1-rank step timed in THIS run (rank 0 alone; the others wait): the denominator of weak_speedup_vs_n1
    n1 = None
    if world > 1:
        if rank == 0:
            n1 = _run_one(None, 0, 1, 1024, steps, warmup, e2e=False)
        _lib.check(lib.wf2cu_sync())
        dist.barrier()
    detail["weak"] = _run_one(dist, rank, world, 1024, steps, warmup, exchange=exchange)                 # 1024 per GPU
    if world == 1:
        n1 = detail["weak"]
    if world > 1 and 1024 % world == 0:
        detail["strong"] = _run_one(dist, rank, world, 1024 // world, steps, warmup, exchange=exchange)  # 1024 global
    if world > 1 and exchange == "p2p":
        detail["weak_nccl"] = _run_one(dist, rank, world, 1024, steps, warmup, exchange="nccl")          # the library baseline
    check = None
    if world > 1:
        import torch
        n1_sps = torch.tensor([n1["samples_per_s"] if rank == 0 else 0.0], device="cuda", dtype=torch.float64)
        dist.broadcast(n1_sps, src=0)
        n1_sps = float(n1_sps.item())
        check = {"replicas_identical": all(detail[k]["replicas_identical"] for k in ("weak", "strong", "weak_nccl") if k in detail),
                 "timed_out": any(detail[k]["p2p_timed_out"] for k in ("weak", "strong") if k in detail)}
        if exchange == "p2p":
            check["kernel_ulp_vs_emulation"] = _p2p_kernel_check(dist, rank, world)
            check["bitwise_vs_emulation"] = check["kernel_ulp_vs_emulation"] == 0
            a, b = np.array(detail["weak"]["loss_first20"]), np.array(detail["weak_nccl"]["loss_first20"])
            k = min(len(a), len(b))
            check["loss_first20_p2p_vs_nccl"] = float(np.max(np.abs(a[:k] - b[:k]) / np.abs(b[:k]))) if k else None
        check["loss_first5_vs_oracle"] = _oracle_check(dist, rank, world, exchange, host_losses) if host_losses else None
        check["ok"] = bool(check["replicas_identical"] and not check["timed_out"] and check.get("bitwise_vs_emulation", True)
                           and (check["loss_first5_vs_oracle"] is None or check["loss_first5_vs_oracle"] <= 1e-4))
        _lib.check(lib.wf2cu_comm_destroy())
    else:
        n1_sps = n1["samples_per_s"]
    cpu = None
    if cpu_fn is not None and rank == 0 and world == 1:
        try:
            cpu = cpu_fn()
        except Exception as e:                                   # side information: may never break the bench line
            cpu = {"error": repr(e)}
    detail["n1_same_run"] = n1
    detail["p2p_check"] = check
    detail["cpu_baseline"] = cpu

    def brief(d):
        return None if d is None else {"sps": _r(d["samples_per_s"], 0), "us": _r(d["ms_per_step"] * 1e3, 1),
                                       "e2e_sps": _r(d["e2e_samples_per_s"], 0), "loss": _r(d["loss_after"], 2)}
    compact = {"unit": "samples/s", "n_gpus": world, "exchange": exchange if world > 1 else None,
               "weak": brief(detail["weak"]), "strong": brief(detail.get("strong")), "weak_nccl": brief(detail.get("weak_nccl")),
               "n1_same_run": {"sps": _r(n1_sps, 0)},
               "weak_speedup_vs_n1": _r(detail["weak"]["samples_per_s"] / n1_sps, 3),
               "strong_speedup_vs_n1": _r(detail["strong"]["samples_per_s"] / n1_sps, 3) if "strong" in detail else None,
               "nccl_speedup_vs_n1": _r(detail["weak_nccl"]["samples_per_s"] / n1_sps, 3) if "weak_nccl" in detail else None,
               "launches_per_step": detail["weak"]["launches_per_step"],
               "p2p_check": {k: (round(v, 9) if isinstance(v, float) else v) for k, v in check.items()} if check else None,
               "cpu_baseline": {"sps": _r(cpu["value"], 0), "cores": cpu["cores"], "kind": cpu["kind"]} if cpu and "value" in cpu else None}
    return compact, detail


def run_cnn(wf, steps: int = 100, warmup: int = 10) -> dict:
    """§8 f-3: the MNIST-CNN training step (examples/mnist/cnn.lisp:5-26; Conv2D 1->4 3x3, relu, MaxPool 2,
    Conv2D 4->16 5x5, relu, MaxPool 2, LinearLayer 256->10, softmax-cross-entropy (!sum), Adam) on one GPU, at
    the example's batch of 100 and at 1024."""
    res = {"model": "MNIST-CNN (examples/mnist/cnn.lisp): Conv2D(1,4,3x3) relu MaxPool2D(2) Conv2D(4,16,5x5) relu "
                    "MaxPool2D(2) LinearLayer(256,10), softmax-cross-entropy (!sum), Adam lr 1e-3, fp32",
           "data": "synthetic U(0,1) images (N 1 28 28), one-hot labels"}
    for b in (100, 1024):
        d = _run_one(None, 0, 1, b, steps, warmup, workload="cnn")
        d.pop("loss_first20", None)
        res[f"batch{b}"] = d
    return res
