#!/usr/bin/env python
"""bench.py — headline benchmark of the CUDATensor hot path (BASELINE.json metric
"fused elementwise/reduce HBM GB/s; MLP train samples/s at 1/2/4/8 B200").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--size 4096]

A "step" is one pass of the C2 sweep over fp32 (4096 4096) tensors, every case one
call through the C ABI of libwaffe2cuda.so:
    add, mul x bcast(4096,1), mul x bcast(1,4096), sum axis 1 / 0 / t, softmax (axis 1, mean-shift),
    exp, sin, tanh, in-place A+=B, add x (!t y)
`value` = algorithmic bytes of the sweep / device time per step (GB/s), inputs resident in HBM.
Inputs rotate through NSETS buffer sets (> 2x the 126 MB L2) so every launch reads HBM-cold data.
At N > 1 each rank runs an independent replica of the sweep ("replicas only") and the
value is the aggregate; the batch-sharded MLP training step (the one sharded path) is reported in
the `mlp` object of the same line when --mlp is on (default).

`--impl reference` times the reference's own CPU kernels (oracle/_ref, driven by the restated Lisp
loop nest) on the host cores for the same sweep.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "fused_elementwise_reduce_hbm_gbps"
UNIT = "GB/s"
NSETS = 4


# ------------------------------------------------------------------------------- sweep definition
def sweep_cases(n: int):
    """(name, algorithmic bytes) per case — SURVEY §8(d) C2 (i)-(viii) for an (n n) fp32 tensor."""
    t = n * n * 4
    row = n * 4
    return [
        ("add", 3 * t), ("mul_bcast_axis1", 2 * t + row), ("mul_bcast_axis0", 2 * t + row),
        ("sum_axis1", t + row), ("sum_axis0", t + row), ("sum_all", t + 4),
        ("softmax", 2 * t), ("exp", 2 * t), ("sin", 2 * t), ("tanh", 2 * t),
        ("add_inplace", 3 * t), ("add_transposed", 3 * t),
    ]


class GpuSweep:
    """The sweep on CUDATensor through the public Python host API (one C-ABI call per case)."""

    def __init__(self, wf, n: int, seed: int = 0):
        self.wf, self.n = wf, n
        rng = np.random.default_rng(seed)
        self.hx = rng.uniform(-1, 1, size=(n, n)).astype(np.float32)
        self.hy = rng.uniform(-1, 1, size=(n, n)).astype(np.float32)
        self.hcol = rng.uniform(-1, 1, size=(n, 1)).astype(np.float32)
        self.hrow = rng.uniform(-1, 1, size=(1, n)).astype(np.float32)
        self.sets = []
        for _ in range(NSETS):
            s = dict(x=wf.from_numpy(self.hx), y=wf.from_numpy(self.hy), out=wf.make_tensor((n, n)),
                     col=wf.from_numpy(self.hcol), row=wf.from_numpy(self.hrow),
                     s1=wf.make_tensor((n, 1)), s0=wf.make_tensor((1, n)), sa=wf.make_tensor((1, 1)))
            s["colb"] = wf.view(s["col"], True, [":broadcast", n])
            s["rowb"] = wf.view(s["row"], [":broadcast", n], True)
            s["yt"] = wf.transpose(s["y"])
            self.sets.append(s)
        self.cursor = 0
        self.cases = sweep_cases(n)

    def run_case(self, name: str, s: dict):
        o = self.wf.ops
        if name == "add": o.add(s["x"], s["y"], out=s["out"])
        elif name == "mul_bcast_axis1": o.mul(s["x"], s["colb"], out=s["out"])
        elif name == "mul_bcast_axis0": o.mul(s["x"], s["rowb"], out=s["out"])
        elif name == "sum_axis1": o.sum(s["x"], axis=1, out=s["s1"])
        elif name == "sum_axis0": o.sum(s["x"], axis=0, out=s["s0"])
        elif name == "sum_all": o.sum(s["x"], axis=True, out=s["sa"])
        elif name == "softmax": o.softmax(s["x"], out=s["out"])
        elif name in ("exp", "sin", "tanh"): o.unary(name, s["x"], out=s["out"])
        elif name == "add_inplace": o.add_(s["out"], s["y"])
        elif name == "add_transposed": o.add(s["x"], s["yt"], out=s["out"])
        else: raise KeyError(name)

    def step(self):
        for name, _ in self.cases:
            self.cursor = (self.cursor + 1) % NSETS
            self.run_case(name, self.sets[self.cursor])

    def bytes_per_step(self) -> int:
        return sum(b for _, b in self.cases)


class CpuSweep:
    """The same sweep on the oracle (reference C kernels driven by the restated loop nest)."""

    def __init__(self, n: int, seed: int = 0):
        from oracle import kernels, loopnest
        self.L, self.n = loopnest, n
        self.kind = kernels.load().kind
        rng = np.random.default_rng(seed)
        L = loopnest
        self.x = L.make_tensor((n, n), data=rng.uniform(-1, 1, size=(n, n)).astype(np.float32))
        self.y = L.make_tensor((n, n), data=rng.uniform(-1, 1, size=(n, n)).astype(np.float32))
        self.col = L.make_tensor((n, 1), data=rng.uniform(-1, 1, size=(n, 1)).astype(np.float32))
        self.row = L.make_tensor((1, n), data=rng.uniform(-1, 1, size=(1, n)).astype(np.float32))
        self.colb = L.view(self.col, True, [":broadcast", n])
        self.rowb = L.view(self.row, [":broadcast", n], True)
        self.yt = L.transpose(self.y)
        self.acc = L.copy(self.x)
        self.cases = sweep_cases(n)

    def run_case(self, name: str):
        L = self.L
        if name == "add": return L.binary("add", self.x, self.y)
        if name == "mul_bcast_axis1": return L.binary("mul", self.x, self.colb)
        if name == "mul_bcast_axis0": return L.binary("mul", self.x, self.rowb)
        if name == "sum_axis1": return L.sum_(self.x, axis=1)
        if name == "sum_axis0": return L.sum_(self.x, axis=0)
        if name == "sum_all": return L.sum_(self.x, axis=True)
        if name == "softmax": return L.softmax(self.x)
        if name in ("exp", "sin", "tanh"): return L.unary(name, self.x)
        if name == "add_inplace": return L.arith_("add", self.acc, self.y)
        if name == "add_transposed": return L.binary("add", self.x, self.yt)
        raise KeyError(name)

    def step(self, threads: int = 1):
        if threads <= 1:
            for name, _ in self.cases:
                self.run_case(name)
            return
        import concurrent.futures as cf
        with cf.ThreadPoolExecutor(max_workers=threads) as ex:   # ctypes calls release the GIL
            list(ex.map(self.run_case, [c for c, _ in self.cases]))

    def bytes_per_step(self) -> int:
        return sum(b for _, b in self.cases)


# ------------------------------------------------------------------------------- helpers
class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index, self.samples, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                r = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                    "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if r.returncode == 0 and r.stdout.strip():
                    self.samples.append([f.strip() for f in r.stdout.strip().split(",")])
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for s in self.samples for i in range(4) if len(s) > 2 + i and s[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(key):
    """dram read + write bytes per launch of `key`, from the committed ncu --set full capture
    (profiles/ncu_traffic.json; written by hand from profiles/r*_ncu_full_*.md).  None if never captured."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        return json.load(open(p)).get(key)
    except (OSError, ValueError):
        return None


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))



def gemm_sweep(wf, sizes=(1024, 2048, 4096, 8192)):
    """C3: `!matmul` (N N)x(N N) fp32 through wf2cu_gemm (auto mode -> tcgen05 3xTF32).  flops = 2N^3
    (algorithmic); the tensor pipe executes 3 TF32 passes, so pipe-level = 3x.  Peak: TF32 dense =
    1/2 x the measured bf16 figure of MEASURED_PEAKS.json (no TF32 figure is driver-measured);
    cuBLAS TF32 on this pool measured 715 TFLOP/s (profiles/r01_cublas_gemm_peaks.log)."""
    import ctypes as C
    lib = wf._lib.load()
    chk = wf._lib.check
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    bf16 = float(json.load(open(p))["bf16_tflops"]) if os.path.exists(p) else 1590.0
    tf32_peak = bf16 / 2
    rng = np.random.default_rng(3)
    cases = []
    for N in sizes:
        A = wf.from_numpy(rng.uniform(-1, 1, size=(N, N)).astype(np.float32))
        B = wf.from_numpy(rng.uniform(-1, 1, size=(N, N)).astype(np.float32))
        o = wf.make_tensor((N, N))
        for _ in range(3):
            wf.ops.matmul(A, B, out=o)
        reps = 10 if N <= 4096 else 5
        best = None
        for _ in range(2):
            chk(lib.wf2cu_timer_start())
            for _ in range(reps):
                wf.ops.matmul(A, B, out=o)
            ms = C.c_float()
            chk(lib.wf2cu_timer_stop(C.byref(ms)))
            best = ms.value / reps if best is None else min(best, ms.value / reps)
        tf = 2.0 * N ** 3 / (best * 1e-3) / 1e12
        cases.append({"N": N, "ms": best, "tflops": tf, "frac_of_tf32_peak": tf / tf32_peak, "pipe_tflops": 3 * tf,
                      "pipe_frac": 3 * tf / tf32_peak})
        del A, B, o
    top = cases[-1]
    return {"workload": "c3_matmul_NxN_f32 (split planes + tcgen05 3xTF32, all launches timed; N <= 1024 the one-launch fused kernel, N = 2048 128x256 tiles, N >= 4096 256x256 tiles)", "cases": cases,
            "roofline": {"bound": "tensor", "kernel": "gemm_tc256_kernel (tcgen05.mma kind::tf32, 3 passes, 256x256 tiles)", "achieved": top["tflops"],
                         "peak": tf32_peak, "unit": "TFLOP/s", "frac": top["frac_of_tf32_peak"], "pipe_frac": top["pipe_frac"],
                         "peak_source": "0.5 x MEASURED_PEAKS.json bf16_tflops (TF32 dense); of measured", "N": top["N"],
                         "traffic": None, "traffic_at_N4096": ncu_traffic("gemm_tc_kernel_N4096")}}


def cpu_sgemm_baseline(n=2048, reps=3):
    """the reference's MatmulNode leaf is BLAS `sgemm_` (`source/backends/cpu/matrix-ops.lisp:58-79`); the stand-in
    here is the OpenBLAS behind numpy (the reference pins no BLAS version).  Timed with all host threads and with one."""
    rng = np.random.default_rng(5)
    a = rng.uniform(-1, 1, size=(n, n)).astype(np.float32)
    b = rng.uniform(-1, 1, size=(n, n)).astype(np.float32)

    def timed():
        a @ b
        best = None
        for _ in range(reps):
            t0 = time.perf_counter()
            a @ b
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        return 2.0 * n ** 3 / best / 1e9
    out = {"N": n, "kind": "OpenBLAS sgemm through numpy (stand-in for the reference's unpinned system BLAS)", "host_cores": os.cpu_count()}
    out["gflops_all_threads"] = timed()
    try:
        from threadpoolctl import threadpool_limits
        with threadpool_limits(limits=1):
            out["gflops_1_thread"] = timed()
    except Exception as e:                                  # threadpoolctl missing: report only the all-thread figure
        out["gflops_1_thread"] = None
        out["note"] = repr(e)
    return out


def gemm_numerics(wf, n=1024):
    """max|C - C64| / max|C64| at N = n for the device path and for OpenBLAS fp32 beside it."""
    rng = np.random.default_rng(9)
    a = rng.uniform(-1, 1, size=(n, n)).astype(np.float32)
    b = rng.uniform(-1, 1, size=(n, n)).astype(np.float32)
    c64 = a.astype(np.float64) @ b.astype(np.float64)
    ours = wf.ops.matmul(wf.from_numpy(a), wf.from_numpy(b)).to_numpy().astype(np.float64)
    blas = (a @ b).astype(np.float64)
    den = float(np.abs(c64).max())
    return {"N": n, "ours_max_rel_err": float(np.abs(ours - c64).max()) / den,
            "openblas_fp32_max_rel_err": float(np.abs(blas - c64).max()) / den, "tolerance": 1e-4}


def gpt2_forward_bench(wf, batch=8, seq=1024, reps=3):
    """C5: examples/gpt-2 GPT-2 small forward (hidden states, as `(call model prev)` returns them),
    random-init weights, fp32, B=8, N=1024, semantics as the example writes them.  The launch sequence
    runs under a StaticPool (the reference's static allocation) and is replayed as one CUDA graph."""
    import ctypes as C
    lib = wf._lib.load()
    chk = wf._lib.check
    g = importlib.import_module("cl-waffe2_b200.gpt2")
    cfg = dict(g.CONFIG_SMALL)
    model = g.GPT2(cfg, seed=0)
    rng = np.random.default_rng(11)
    ctx = wf.from_numpy(rng.integers(0, cfg["n_vocab"], size=(batch, seq)).astype(np.float32))
    pool = wf.StaticPool()
    l0 = C.c_int64(); l1 = C.c_int64()
    for _ in range(2):                       # 1st run allocates the plan + GEMM workspace, 2nd checks it is static
        chk(lib.wf2cu_launch_count(C.byref(l0)))
        with pool:
            out = model(ctx)
        chk(lib.wf2cu_launch_count(C.byref(l1)))
    chk(lib.wf2cu_sync())
    pool.frozen = True
    ex = C.c_void_p()
    chk(lib.wf2cu_graph_begin())
    try:
        with pool:
            out = model(ctx)
    finally:
        rc = lib.wf2cu_graph_end(C.byref(ex))
    chk(rc)
    chk(lib.wf2cu_graph_launch(ex))
    best = None
    for _ in range(reps):
        chk(lib.wf2cu_timer_start())
        chk(lib.wf2cu_graph_launch(ex))
        ms = C.c_float()
        chk(lib.wf2cu_timer_stop(C.byref(ms)))
        best = ms.value if best is None else min(best, ms.value)
    h = out.to_numpy()
    chk(lib.wf2cu_graph_destroy(ex))
    # matmul flops of one forward: per layer qkv + attn-proj + fc + fc-proj, plus QK^T and PV
    E, L, H = cfg["n_emb"], cfg["n_layer"], cfg["n_head"]
    T = batch * seq
    flops = L * (2 * T * E * 3 * E + 2 * T * E * E + 2 * 2 * T * E * 4 * E + 2 * 2 * batch * H * seq * seq * (E // H))
    return {"workload": f"gpt2_small_forward_b{batch}_n{seq}_f32 (hidden states, random-init weights)", "ms_per_forward": best,
            "tokens_per_s": T / (best * 1e-3), "launches_per_forward": int(l1.value - l0.value),
            "matmul_tflops": flops / (best * 1e-3) / 1e12, "pool_bytes": pool.nbytes(),
            "finite": bool(np.isfinite(h).all()), "hidden_absmax": float(np.abs(h).max())}

# ------------------------------------------------------------------------------- reference arm
def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    n = args.ref_size
    cores = os.cpu_count() or 1
    threads = min(cores, len(sweep_cases(n)))
    sw = CpuSweep(n)
    for _ in range(args.warmup):
        sw.step(threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sw.step(threads)
    dt = (time.perf_counter() - t0) / args.steps
    val = sw.bytes_per_step() / dt / 1e9
    sample = f"full C2 sweep on ({n} {n}) fp32, {len(sw.cases)} cases run concurrently on {threads} host threads " \
             f"(each reference kernel is single-threaded, as in cl-waffe2: *num-cores*=1)"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"c2_elementwise_reduce_sweep_{n}x{n}_f32", "cases": [c for c, _ in sw.cases],
                   "bytes_per_step": sw.bytes_per_step()},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": sw.kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------- our arm
def stage(msg: str):
    """Progress marker on stderr (WF2_BENCH_TRACE=1): tells where a multi-rank run stopped."""
    if os.environ.get("WF2_BENCH_TRACE"):
        print(f"[bench rank {os.environ.get('RANK', '0')} t={time.time() % 1000:.1f}] {msg}", file=sys.stderr, flush=True)


def run_ours(args):
    rank, local_rank, world = dist_env()
    if os.environ.get("WF2_BENCH_TRACE"):
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ.get("WF2_BENCH_TRACE_AFTER", "90")), exit=False, file=sys.stderr)
    if world != args.gpus and world > 1:
        args.gpus = world
    wf = importlib.import_module("cl-waffe2_b200")
    wf.init(local_rank)
    lib = wf._lib.load()
    chk = wf._lib.check
    import ctypes as C

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        chk(lib.wf2cu_sync())
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    stage("process group ready")
    n = args.size
    sw = GpuSweep(wf, n)
    stage("sweep buffers uploaded")
    launches0 = C.c_int64()

    def capture(fn):
        """Record fn()'s launches into a CUDA graph (the static iseq of a compiled model)."""
        ex = C.c_void_p()
        chk(lib.wf2cu_graph_begin())
        fn()
        chk(lib.wf2cu_graph_end(C.byref(ex)))
        return ex

    # ---- timed region: K steps, device time on the library stream.  One step (12 C-ABI calls) is
    # captured once into a CUDA graph and replayed, so the GPU is not waiting on the Python host.
    for _ in range(max(args.warmup, 3)):
        sw.step()
    sw.cursor = 0
    step_graph = capture(sw.step)
    for _ in range(max(args.warmup, 3)):
        chk(lib.wf2cu_graph_launch(step_graph))
    barrier()
    with ClockSampler(local_rank) as clocks:
        chk(lib.wf2cu_launch_count(C.byref(launches0)))
        chk(lib.wf2cu_timer_start())
        for _ in range(args.steps):
            chk(lib.wf2cu_graph_launch(step_graph))
        ms = C.c_float()
        chk(lib.wf2cu_timer_stop(C.byref(ms)))
        launches1 = C.c_int64()
        chk(lib.wf2cu_launch_count(C.byref(launches1)))
        barrier()
        # keep the sampler alive long enough for a few samples on very short runs
        t_end = time.time() + 0.5
        while time.time() < t_end:
            for _ in range(20):
                chk(lib.wf2cu_graph_launch(step_graph))
            chk(lib.wf2cu_sync())
    stage("sweep timed")
    total_ms = max_over_ranks(float(ms.value))
    ms_per_step = total_ms / args.steps
    bytes_step = sw.bytes_per_step()
    value = world * bytes_step / (ms_per_step * 1e-3) / 1e9

    # the same K steps issued call by call from Python (no graph), for reference
    barrier()
    chk(lib.wf2cu_timer_start())
    for _ in range(args.steps):
        sw.step()
    t = C.c_float()
    chk(lib.wf2cu_timer_stop(C.byref(t)))
    eager_ms = t.value / args.steps

    # ---- per-case device time: `reps` launches of one case (rotating buffer sets) in one graph
    per_case = {}
    reps = 20
    for name, nbytes in sw.cases:
        def many(name=name):
            for _ in range(reps):
                sw.cursor = (sw.cursor + 1) % NSETS
                sw.run_case(name, sw.sets[sw.cursor])
        many()
        g = capture(many)
        chk(lib.wf2cu_graph_launch(g))
        best = None
        for _ in range(3):
            chk(lib.wf2cu_timer_start())
            chk(lib.wf2cu_graph_launch(g))
            t = C.c_float()
            chk(lib.wf2cu_timer_stop(C.byref(t)))
            best = t.value if best is None else min(best, t.value)
        chk(lib.wf2cu_graph_destroy(g))
        per_case[name] = {"ms": best / reps, "bytes": nbytes, "gbps": nbytes / (best / reps * 1e-3) / 1e9}
    peak, peak_src = measured_peaks()
    for c in per_case.values():
        c["frac"] = c["gbps"] / peak
    dom = "add"   # ew_vec_kernel<2, add>: the largest share of the step's bytes (3 cases use it)
    roofline = {"bound": "hbm", "kernel": "ew_vec_kernel<2,FBinary<ADD>> (case add)", "achieved": per_case[dom]["gbps"],
                "peak": peak, "unit": "GB/s", "frac": per_case[dom]["gbps"] / peak, "peak_source": peak_src,
                "traffic": ncu_traffic("ew_vec_kernel<2,FBinary<ADD>>"), "traffic_source": "profiles/ncu_traffic.json (ncu --set full capture, dram read+write per launch)",
                "algorithmic_bytes_per_launch": per_case[dom]["bytes"]}

    # ---- e2e: host buffers in, host results out, copies inside the timed region
    hx = C.c_void_p(); hy = C.c_void_p(); hres = C.c_void_p()
    nb = n * n * 4
    chk(lib.wf2cu_host_alloc(nb, C.byref(hx))); chk(lib.wf2cu_host_alloc(nb, C.byref(hy)))
    d2h_bytes = 2 * n * 4 + 4
    chk(lib.wf2cu_host_alloc(d2h_bytes, C.byref(hres)))
    C.memmove(hx.value, sw.hx.ctypes.data, nb); C.memmove(hy.value, sw.hy.ctypes.data, nb)

    def upload(s):
        chk(lib.wf2cu_h2d(s["x"].vec.ptr, 0, hx, nb))
        chk(lib.wf2cu_h2d(s["y"].vec.ptr, 0, hy, nb))

    def e2e_run(steps):
        """`steps` sweeps, each on inputs copied from pinned host memory inside the timed region.  The copy for
        step i+1 goes to a side stream (wf2cu_branch_*) into the next buffer set while step i computes; step
        0's copy is exposed.  Exactly `steps` uploads, `steps` sweeps and `steps` result read-backs."""
        cur = sw.sets[sw.cursor]
        upload(cur)
        for i in range(steps):
            nxt = sw.sets[(sw.cursor + 1) % NSETS]
            if i + 1 < steps:
                chk(lib.wf2cu_branch_begin(1))
                try:
                    upload(nxt)
                finally:
                    chk(lib.wf2cu_branch_end(1))
            for name, _ in sw.cases:
                sw.run_case(name, cur)
            chk(lib.wf2cu_d2h(cur["s1"].vec.ptr, 0, hres.value, n * 4))
            chk(lib.wf2cu_d2h(cur["s0"].vec.ptr, 0, hres.value + n * 4, n * 4))
            chk(lib.wf2cu_d2h(cur["sa"].vec.ptr, 0, hres.value + 2 * n * 4, 4))
            if i + 1 < steps:
                chk(lib.wf2cu_branch_join(1))
            sw.cursor = (sw.cursor + 1) % NSETS
            cur = nxt

    e2e_run(3)
    barrier()
    e2e_steps = max(3, min(args.steps, 20))
    chk(lib.wf2cu_timer_start())
    e2e_run(e2e_steps)
    t = C.c_float()
    chk(lib.wf2cu_timer_stop(C.byref(t)))
    e2e_ms = max_over_ranks(float(t.value)) / e2e_steps
    e2e = {"value": world * bytes_step / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e2e_ms,
           "h2d_bytes_per_step": 2 * nb, "d2h_bytes_per_step": d2h_bytes,
           "api": "cl-waffe2_b200.ops (one libwaffe2cuda C-ABI call per node) + wf2cu_h2d/wf2cu_d2h from pinned host memory; "
                  "the upload of step i+1 runs on a side stream (wf2cu_branch_*) under step i's kernels"}

    stage("e2e timed")
    # ---- MLP training step (the sharded path)
    mlp = None
    if args.mlp:
        try:
            mlp_mod = importlib.import_module("cl-waffe2_b200.mlp_bench")
            mlp = mlp_mod.run(wf, dist, rank, world, steps=args.mlp_steps, warmup=args.mlp_warmup, exchange=args.exchange)
        except ModuleNotFoundError:
            mlp = {"unavailable": "MLP training step not built yet"}

    stage("mlp done")
    # ---- MatmulNode sweep (C3), single GPU only
    gemm = None
    if args.gemm and world == 1:
        gemm = gemm_sweep(wf)
        for key, fn in (("numerics", lambda: gemm_numerics(wf)), ("cpu_baseline", cpu_sgemm_baseline)):
            if key == "cpu_baseline" and (rank != 0 or args.no_cpu):
                continue
            try:                                             # side information: may never break the bench line
                gemm[key] = fn()
            except Exception as e:
                gemm[key] = {"error": repr(e)}

    gpt2 = None
    if args.gpt2 and world == 1:
        stage("gpt2")
        try:
            gpt2 = gpt2_forward_bench(wf)
        except Exception as e:                       # reported, never silently dropped
            gpt2 = {"error": repr(e)}

    cnn = None
    if args.cnn and world == 1:
        stage("cnn")
        try:
            cnn = importlib.import_module("cl-waffe2_b200.mlp_bench").run_cnn(wf)
        except Exception as e:
            cnn = {"error": repr(e)}

    # ---- CPU baseline (rank 0, N=1): the oracle on one host core, bounded sample
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cn = args.cpu_size
        cs = CpuSweep(cn)
        cs.step(1)
        t0 = time.perf_counter()
        reps = 0
        while reps < 1 or (time.perf_counter() - t0 < 8 and reps < 5):
            cs.step(1)
            reps += 1
        dt = (time.perf_counter() - t0) / reps
        cpu = {"value": cs.bytes_per_step() / dt / 1e9, "unit": UNIT, "cores": 1, "kind": cs.kind,
               "sample": f"the same 12-case sweep on ({cn} {cn}) fp32, {reps} passes after 1 warm-up, 1 thread "
                         f"(reference default *num-cores*=1); host has {os.cpu_count()} cores",
               "ms_per_step": dt * 1e3}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "eager_ms_per_step": eager_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"c2_elementwise_reduce_sweep_{n}x{n}_f32", "cases": [c for c, _ in sw.cases],
                       "bytes_per_step": bytes_step, "l2": f"inputs rotate through {NSETS} buffer sets "
                       f"({NSETS * 3 * nb >> 20} MiB > 2x L2), no flush kernel in the timed region",
                       "launch": "one step captured into a CUDA graph and replayed",
                       "parallelism": "replicas only" if world > 1 else "single GPU"},
            "roofline": roofline, "cases": per_case, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": int(launches1.value - launches0.value), "clocks": clocks.summary(), "mlp": mlp, "gemm": gemm, "gpt2": gpt2,
            "cnn": cnn,
        }
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=4096)
    ap.add_argument("--ref-size", type=int, default=4096)
    ap.add_argument("--cpu-size", type=int, default=4096)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--mlp", type=int, default=1)
    ap.add_argument("--gemm", type=int, default=1)
    ap.add_argument("--gpt2", type=int, default=1)
    ap.add_argument("--cnn", type=int, default=1)
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="N>1 gradient exchange: p2p = fused peer-memory allreduce+Adam kernel, nccl = ncclAllReduce + Adam")
    ap.add_argument("--mlp-steps", type=int, default=200)
    ap.add_argument("--mlp-warmup", type=int, default=20)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
