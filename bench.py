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
NOMINAL_HBM_GBS = 8000.0          # north_star's "~8 TB/s" denominator, reported beside the measured one


def workload_config(n: int, cases):
    """`config` of the JSON line — byte-identical in both arms (the driver compares them)."""
    nb = n * n * 4
    return {"workload": f"c2_elementwise_reduce_sweep_{n}x{n}_f32", "cases": [c for c, _ in cases],
            "bytes_per_step": sum(b for _, b in cases),
            "l2": f"inputs larger than L2: every case streams 64 MiB tensors and the GPU arm rotates through {NSETS} buffer sets "
                  f"({NSETS * 3 * nb >> 20} MiB > 2x the 126 MB L2); no flush kernel in the timed region"}


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

    def step(self, branches: int = 1):
        """One pass over the 12 cases.  branches > 1: the cases are independent nodes (each reads and writes its own buffer
        set), so a dataflow executor need not serialise them: they alternate between `branches` side branches of the compiled
        sequence (wf2cu_branch_begin/end/join, `with-cuda-branch` on the Lisp side) and the launch boundaries of one chain
        overlap the streaming of the other."""
        lib = self.wf._lib.load()
        chk = self.wf._lib.check
        for i, (name, _) in enumerate(self.cases):
            self.cursor = (self.cursor + 1) % NSETS
            if branches > 1:
                b = 1 + i % branches
                chk(lib.wf2cu_branch_begin(b))
                try:
                    self.run_case(name, self.sets[self.cursor])
                finally:
                    chk(lib.wf2cu_branch_end(b))
            else:
                self.run_case(name, self.sets[self.cursor])
        if branches > 1:
            for b in range(1, branches + 1):
                chk(lib.wf2cu_branch_join(b))

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



GEMM_PATHS = {0: "none", 1: "simt", 2: "simt_splitk", 3: "simt_skinny", 4: "tc_fused", 5: "tc_128x256", 6: "tc_256x256", 7: "tc_256x192", 8: "tc_raw", 9: "tc_2sm"}


def _gemm_rows_check(rng, A, B, got, ta, tb, nrows=64):
    """max|C - C64| / max|C64| over `nrows` sampled rows of the output of the TIMED launches, and OpenBLAS fp32's own figure."""
    N = A.shape[0]
    rows = np.unique(np.concatenate([[0, N // 2, N - 1], rng.integers(0, N, size=nrows)]))
    opA = (A.T if ta else A)[rows]
    opB = B.T if tb else B
    truth = opA.astype(np.float64) @ opB.astype(np.float64)
    blas = (np.ascontiguousarray(opA) @ np.ascontiguousarray(opB)).astype(np.float64)
    den = float(np.abs(truth).max())
    return float(np.abs(got[rows].astype(np.float64) - truth).max()) / den, float(np.abs(blas - truth).max()) / den


def gemm_sweep(wf, sizes=(256, 512, 1024, 2048, 4096, 8192), check=True):
    """C3: `!matmul` (N N)x(N N) fp32 through wf2cu_gemm under AUTO dispatch (N = 256..8192 and all four `trans`
    combos at N = 4096).  flops = 2N^3 (algorithmic).  The tensor pipe executes more: the plane-fed kernels run one TF32 MMA
    plus two bf16 correction MMAs per k-step (bf16 runs at twice the TF32 rate: 2 TF32-equivalent units, `pipe_units`), the
    raw-operand and small-problem kernels three TF32 MMAs (3 units); pipe-level = units x.  Peak: TF32 dense = 1/2 x the
    measured bf16 figure of MEASURED_PEAKS.json (no TF32 figure is driver-measured); cuBLAS TF32 is measured in the same run
    (`cublas_tf32`).  `err` is computed on the output buffer of the timed launches."""
    import ctypes as C
    lib = wf._lib.load()
    chk = wf._lib.check
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    bf16 = float(json.load(open(p))["bf16_tflops"]) if os.path.exists(p) else 1590.0
    tf32_peak = bf16 / 2
    rng = np.random.default_rng(3)
    cases = []
    plan = [(N, False, False) for N in sizes] + [(4096, True, False), (4096, False, True), (4096, True, True)]
    for N, ta, tb in plan:
        hA = rng.uniform(-1, 1, size=(N, N)).astype(np.float32)
        hB = rng.uniform(-1, 1, size=(N, N)).astype(np.float32)
        A, B = wf.from_numpy(hA), wf.from_numpy(hB)
        o = wf.make_tensor((N, N))
        a_op = wf.ops.t(A) if ta else A
        b_op = wf.ops.t(B) if tb else B
        for _ in range(3):
            wf.ops.matmul(a_op, b_op, out=o)
        path = C.c_int(0)
        chk(lib.wf2cu_gemm_last_path(C.byref(path)))
        reps = 20 if N <= 2048 else (10 if N <= 4096 else 5)
        best = None
        for _ in range(2):
            chk(lib.wf2cu_timer_start())
            for _ in range(reps):
                wf.ops.matmul(a_op, b_op, out=o)
            ms = C.c_float()
            chk(lib.wf2cu_timer_stop(C.byref(ms)))
            best = ms.value / reps if best is None else min(best, ms.value / reps)
        tf = 2.0 * N ** 3 / (best * 1e-3) / 1e12
        units = 2 if path.value in (5, 6, 7, 9) else 3       # planes kernels (single CTA and CTA pair): TF32 + 2 x bf16 corrections; others: 3 x TF32
        c = {"N": N, "trans": f"{'T' if ta else 'N'}{'T' if tb else 'N'}", "path": GEMM_PATHS.get(path.value, str(path.value)), "ms": best,
             "tflops": tf, "frac_of_tf32_peak": tf / tf32_peak, "pipe_units": units, "pipe_tflops": units * tf,
             "pipe_frac": units * tf / tf32_peak}
        if check:
            try:
                c["err"], c["openblas_fp32_err"] = _gemm_rows_check(rng, hA, hB, o.to_numpy(), ta, tb)
                c["err_ok"] = bool(c["err"] <= 1e-4)
            except Exception as e:                              # side information: may never break the bench line
                c["err"] = repr(e)
        cases.append(c)
        del A, B, o
    top = max((c for c in cases if c["trans"] == "NN"), key=lambda c: c["N"])
    return {"workload": "c3_matmul_NxN_f32, auto dispatch, every launch of a MatmulNode timed (operand pre-passes included)", "cases": cases,
            "tolerance": 1e-4,
            "roofline": {"bound": "tensor", "kernel": top["path"], "achieved": top["tflops"],
                         "peak": tf32_peak, "unit": "TFLOP/s", "frac": top["frac_of_tf32_peak"], "pipe_frac": top["pipe_frac"],
                         "pipe_units": top["pipe_units"],
                         "peak_source": "0.5 x MEASURED_PEAKS.json bf16_tflops (TF32 dense); of measured", "N": top["N"],
                         "traffic": ncu_traffic(f"gemm_N{top['N']}"), "traffic_at_N4096": ncu_traffic("gemm_N4096")}}


def cublas_tf32_peak(n=8192, reps=5):
    """cuBLAS TF32 at n^3 through torch (a library GEMM as the second denominator); measured after
    every number of this repo's own kernels has been taken."""
    import torch
    torch.backends.cuda.matmul.allow_tf32 = True
    a = torch.randn(n, n, device="cuda"); b = torch.randn(n, n, device="cuda")
    for _ in range(2):
        a @ b
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = None
    for _ in range(3):
        e0.record()
        for _ in range(reps):
            a @ b
        e1.record(); torch.cuda.synchronize()
        t = e0.elapsed_time(e1) / reps
        best = t if best is None else min(best, t)
    return {"N": n, "ms": best, "tflops": 2.0 * n ** 3 / (best * 1e-3) / 1e12, "how": "torch.matmul fp32, allow_tf32=True (cuBLAS), best of 3 x 5"}


def torch_library_cases(n=4096, reps=20, nsets=4):
    """The reduction / softmax cases of the sweep on the library kernels torch ships (CUB / ATen), timed the same way (`reps`
    launches in a row on rotating buffer sets larger than L2, CUDA events, best of 3): context for the per-case roofline
    fractions — a 64 MiB read-only pass is ~10 us at the measured copy peak, so launch ramp and tail are a visible share of it
    for ANY kernel.  Run after every number of this repo's own kernels has been taken."""
    import torch
    xs = [torch.rand(n, n, device="cuda") * 2 - 1 for _ in range(nsets)]
    ys = [torch.rand(n, n, device="cuda") * 2 - 1 for _ in range(nsets)]
    outs = [torch.empty(n, n, device="cuda") for _ in range(nsets)]
    fns = {"sum_axis1": lambda i: xs[i].sum(1), "sum_axis0": lambda i: xs[i].sum(0), "sum_all": lambda i: xs[i].sum(),
           "softmax": lambda i: torch.softmax(xs[i], 1, out=outs[i]), "add": lambda i: torch.add(xs[i], ys[i], out=outs[i]),
           "exp": lambda i: torch.exp(xs[i], out=outs[i])}
    res = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for name, fn in fns.items():
        for i in range(nsets):
            fn(i)
        torch.cuda.synchronize()
        best = None
        for _ in range(3):
            e0.record()
            for r in range(reps):
                fn(r % nsets)
            e1.record(); torch.cuda.synchronize()
            t = e0.elapsed_time(e1) / reps
            best = t if best is None else min(best, t)
        res[name] = {"ms": best}
    res["how"] = f"torch {torch.__version__} eager ops on ({n} {n}) fp32, {reps} launches in a row over {nsets} buffer sets, CUDA events, best of 3"
    return res


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


def gpt2_forward_bench(wf, batch=8, seq=1024, reps=3, oracle_check=True):
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
    check = None
    if oracle_check:
        # rows of the batch are independent (the heads-axis softmax couples heads, not rows): batch row 0 of the timed
        # B = 8 forward against the numpy restatement of the example in fp64 on the host
        try:
            from oracle import gpt2 as O
            t0 = time.perf_counter()
            h64 = O.forward(np.asarray(ctx.to_numpy()[:1]), g.init_weights(cfg, 0), cfg, np.float64)
            check = {"max_rel_err": float(np.abs(h[:1].astype(np.float64) - h64).max() / np.abs(h64).max()), "tolerance": 1e-3,
                     "against": "oracle/gpt2.py fp64, batch row 0 of the timed forward", "host_seconds": time.perf_counter() - t0}
            check["ok"] = bool(check["max_rel_err"] <= 1e-3)
        except Exception as e:
            check = {"error": repr(e)}
    # matmul flops of one forward: per layer qkv + attn-proj + fc + fc-proj, plus QK^T and PV
    E, L, H = cfg["n_emb"], cfg["n_layer"], cfg["n_head"]
    T = batch * seq
    flops = L * (2 * T * E * 3 * E + 2 * T * E * E + 2 * 2 * T * E * 4 * E + 2 * 2 * batch * H * seq * seq * (E // H))
    return {"workload": f"gpt2_small_forward_b{batch}_n{seq}_f32 (hidden states, random-init weights)", "ms_per_forward": best,
            "tokens_per_s": T / (best * 1e-3), "launches_per_forward": int(l1.value - l0.value),
            "matmul_tflops": flops / (best * 1e-3) / 1e12, "pool_bytes": pool.nbytes(),
            "finite": bool(np.isfinite(h).all()), "hidden_absmax": float(np.abs(h).max()), "check": check,
            "max_rel_err": check.get("max_rel_err") if check else None}

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
        "config": workload_config(n, sw.cases),
        "method": {"launch": "host threads, one reference kernel call per row / per flattened tensor", "parallelism": f"{threads} host threads"},
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


def fused_chain_cases(wf, n: int, peak: float, capture, reps: int = 20):
    """`fused_chains`: the element-wise composites of SURVEY 8 a16 on (n n) fp32 — each ONE launch of a compiled chain kernel
    (csrc/fused.cu) or of the fused optimizer kernel — with their algorithmic bytes, GB/s and fraction of the HBM peak.
    sigmoid `nn/activation.lisp:43`; relu backward mask `:11-26`; MSE difference chain `nn/criterion.lisp:73-79`;
    sin(sin(sin x)) `benchmarks/compose.lisp:21-26`; gelu `nn/activation.lisp:61-66`; SGD on views `optimizers/impls/sgd.lisp:20-26`;
    Adam over a 16 Mi-element arena `optimizers/impls/adam.lisp:45-116`."""
    import ctypes as C
    lib = wf._lib.load()
    chk = wf._lib.check
    P = wf._lib
    imm = wf.ops._imm
    gt = P.P_CMPSEL | (P.CMP["gt"] << 8)
    t = n * n * 4
    rng = np.random.default_rng(21)
    sets = []
    for _ in range(NSETS):                                  # rotate: every launch reads HBM-cold data
        sets.append([wf.from_numpy(rng.uniform(-1, 1, size=(n, n)).astype(np.float32)) for _ in range(3)])
    progs = {
        "sigmoid": (2 * t, lambda v: wf.ops.sigmoid(v[0], out=v[2])),
        "gelu": (2 * t, lambda v: wf.ops.gelu(v[0], out=v[2])),
        "sin_sin_sin": (2 * t, lambda v: wf.ops.fused([(P.P_LOAD, 0, 0, 0), (P.P_UNARY, 0, 0, P.UNARY["sin"]), (P.P_UNARY, 0, 0, P.UNARY["sin"]),
                                                       (P.P_UNARY, 0, 0, P.UNARY["sin"]), (P.P_STORE, 0, 1, 0)], [v[0], v[2]])),
        "relu_backward_mask": (3 * t, lambda v: wf.ops.fused([(P.P_LOAD, 0, 0, 0), (P.P_IMM, 1, 0, imm(0.0)), (gt, 1, 0, 1), (P.P_LOAD, 2, 1, 0),
                                                              (P.P_MUL, 1, 2, 1), (P.P_STORE, 1, 2, 0)], [v[0], v[1], v[2]])),
        "mse_diff": (3 * t, lambda v: wf.ops.fused([(P.P_LOAD, 0, 0, 0), (P.P_LOAD, 1, 1, 0), (P.P_SUB, 0, 0, 1), (P.P_MUL, 0, 0, 0),
                                                    (P.P_STORE, 0, 2, 0)], [v[0], v[1], v[2]])),
        "sgd_on_views": (3 * t, lambda v: wf.ops.fused([(P.P_LOAD, 0, 0, 0), (P.P_LOAD, 1, 1, 0), (P.P_IMM, 2, 0, imm(1e-3)), (P.P_MUL, 1, 1, 2),
                                                        (P.P_SUB, 0, 0, 1), (P.P_STORE, 0, 0, 0)], [v[2], v[1]])),
    }
    out = {}
    cursor = [0]
    path = C.c_int(0)
    for name, (nbytes, fn) in progs.items():
        def many(fn=fn):
            for _ in range(reps):
                cursor[0] = (cursor[0] + 1) % NSETS
                fn(sets[cursor[0]])
        many()
        chk(lib.wf2cu_fused_last_path(C.byref(path)))
        g = capture(many)
        chk(lib.wf2cu_graph_launch(g))
        best = None
        for _ in range(3):
            chk(lib.wf2cu_timer_start())
            chk(lib.wf2cu_graph_launch(g))
            tm = C.c_float()
            chk(lib.wf2cu_timer_stop(C.byref(tm)))
            best = tm.value if best is None else min(best, tm.value)
        chk(lib.wf2cu_graph_destroy(g))
        gbps = nbytes / (best / reps * 1e-3) / 1e9
        out[name] = {"ms": best / reps, "bytes": nbytes, "gbps": gbps, "frac": gbps / peak,
                     "executor": {1: "interpreter", 2: "compiled"}.get(path.value, str(path.value))}
    # Adam over a flat 16 Mi-element arena: reads p, g, m, v and writes p, m, v = 7 x 64 MiB
    cnt = 16 * 1024 * 1024
    arenas = []
    for _ in range(2):
        arenas.append([wf.from_numpy(rng.uniform(-1, 1, size=(cnt,)).astype(np.float32)) for _ in range(4)])
    for ar in arenas:
        ar[3] = wf.from_numpy(rng.uniform(0.1, 1, size=(cnt,)).astype(np.float32))     # v >= 0
    k = [0]

    def adam_many():
        for _ in range(reps):
            k[0] ^= 1
            p_, g_, m_, v_ = arenas[k[0]]
            chk(lib.wf2cu_adam_step(p_.vec.ptr, 0, g_.vec.ptr, 0, m_.vec.ptr, 0, v_.vec.ptr, 0, cnt, 1e-3, None, 0.9, 0.999, 1e-7, 1.0))
    adam_many()
    g = capture(adam_many)
    chk(lib.wf2cu_graph_launch(g))
    best = None
    for _ in range(3):
        chk(lib.wf2cu_timer_start())
        chk(lib.wf2cu_graph_launch(g))
        tm = C.c_float()
        chk(lib.wf2cu_timer_stop(C.byref(tm)))
        best = tm.value if best is None else min(best, tm.value)
    chk(lib.wf2cu_graph_destroy(g))
    nbytes = 7 * cnt * 4
    gbps = nbytes / (best / reps * 1e-3) / 1e9
    out["adam_16Mi_arena"] = {"ms": best / reps, "bytes": nbytes, "gbps": gbps, "frac": gbps / peak, "executor": "adam_kernel"}
    return out


def mlp_host_losses(world: int, per_gpu_batch: int, steps: int, seed: int = 0):
    """CPU checker for the sharded step (bench.py alone may use oracle/): every shard's forward/backward with shared
    parameters, gradients summed over shards x 1/world, one Adam step; the loss is the sum of the shards' `!sum` losses."""
    import importlib as il
    from oracle.mlp import OracleMLP
    nn = il.import_module("cl-waffe2_b200.nn")
    mb = il.import_module("cl-waffe2_b200.mlp_bench")
    ref = OracleMLP(mb.DIMS, "relu", nn.mlp_init_arrays(mb.DIMS, seed=1234), "softmax-cross-entropy-sum", opt="adam", lr=1e-3)
    shards = [mb.shard_data(r, per_gpu_batch, seed) for r in range(world)]
    out = []
    for _ in range(steps):
        total, gW, gb = 0.0, None, None
        for X, Y in shards:
            loss, w, b = ref.forward_backward(X, Y)
            total += loss
            if gW is None:
                gW, gb = w, b
            else:
                for a, c in zip(gW + gb, w + b):
                    a.vec[:] = a.vec + c.vec
        ref.apply(gW, gb, grad_scale=1.0 / world)
        out.append(total)
    return out


def mlp_cpu_baseline(steps: int = 4) -> dict:
    """The reference's own CPU path for the C4 step on the host cores: oracle/mlp.py (its C kernels driven by the restated loop
    nest + OpenBLAS sgemm), batch 1024, a bounded sample.  Context: the reference publishes 3.111 s per epoch of MNIST for
    this model (60 000 samples: 19.3 K samples/s), `Readme.md:140-148`."""
    import importlib as il
    from oracle.mlp import OracleMLP
    nn = il.import_module("cl-waffe2_b200.nn")
    mb = il.import_module("cl-waffe2_b200.mlp_bench")
    X, Y = mb.shard_data(0, 1024, 0)
    ref = OracleMLP(mb.DIMS, "relu", nn.mlp_init_arrays(mb.DIMS, seed=1234), "softmax-cross-entropy-sum", opt="adam", lr=1e-3)
    ref.step(X, Y)
    t0 = time.perf_counter()
    for _ in range(steps):
        ref.step(X, Y)
    dt = (time.perf_counter() - t0) / steps
    return {"value": 1024 / dt, "unit": "samples/s", "ms_per_step": dt * 1e3, "cores": os.cpu_count(), "kind": "reference",
            "sample": f"{steps} Adam steps at batch 1024 after 1 warm-up: reference C kernels (1 thread) + OpenBLAS sgemm (all host threads)"}


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
    warm = max(args.warmup, 3)                               # the timing rules ask for >= 3; what ran is printed as `warmup`

    def capture(fn):
        """Record fn()'s launches into a CUDA graph (the static iseq of a compiled model)."""
        ex = C.c_void_p()
        chk(lib.wf2cu_graph_begin())
        fn()
        chk(lib.wf2cu_graph_end(C.byref(ex)))
        return ex

    # ---- timed region: K steps, device time on the library stream.  One step (12 C-ABI calls) is
    # captured once into a CUDA graph and replayed, so the GPU is not waiting on the Python host.
    nbr = max(1, min(3, args.branches))
    for _ in range(warm):
        sw.step(nbr)
    sw.cursor = 0
    step_graph = capture(lambda: sw.step(nbr))
    for _ in range(warm):
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

    # the same K steps as ONE chain (every case serialised behind the previous one, the reference VM's order): what the step
    # cost before the independent cases were spread over side branches, and the figure rounds 1 and 2a quoted
    single_ms = ms_per_step
    if nbr > 1:
        sw.cursor = 0
        g1 = capture(lambda: sw.step(1))
        for _ in range(warm):
            chk(lib.wf2cu_graph_launch(g1))
        barrier()
        chk(lib.wf2cu_timer_start())
        for _ in range(args.steps):
            chk(lib.wf2cu_graph_launch(g1))
        t1 = C.c_float()
        chk(lib.wf2cu_timer_stop(C.byref(t1)))
        single_ms = max_over_ranks(float(t1.value)) / args.steps
        chk(lib.wf2cu_graph_destroy(g1))
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
    # the dominant kernel = the one with the largest share of the step's time: ew_vec_kernel<2, add> runs 2 of the 12
    # launches (add, add_inplace); the limiting one and the whole-step aggregate are reported beside it
    dom = "add"
    lim = min(per_case, key=lambda k: per_case[k]["frac"])
    agg = bytes_step / (ms_per_step * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "ew_vec_kernel<2,FBinary<ADD>> (case add)", "achieved": per_case[dom]["gbps"],
                "peak": peak, "unit": "GB/s", "frac": per_case[dom]["gbps"] / peak, "peak_source": peak_src,
                "traffic": ncu_traffic("ew_vec_kernel<2,FBinary<ADD>>"),
                "traffic_source": "profiles/ncu_traffic.json (generated by scripts/ncu_traffic.py from the ncu --set full capture: dram read+write per launch)",
                "algorithmic_bytes_per_launch": per_case[dom]["bytes"],
                "aggregate_frac": agg / world / peak if world > 1 else agg / peak, "aggregate_gbps_per_gpu": agg,
                "min_frac": per_case[lim]["frac"], "limiting_kernel": lim,
                "frac_of_nominal_8tbs": per_case[dom]["gbps"] / NOMINAL_HBM_GBS, "aggregate_frac_of_nominal_8tbs": agg / NOMINAL_HBM_GBS,
                "min_frac_of_nominal_8tbs": per_case[lim]["gbps"] / NOMINAL_HBM_GBS}

    fused_chains = None
    if args.fused and world == 1:
        stage("fused chains")
        try:
            fused_chains = fused_chain_cases(wf, n, peak, capture)
        except Exception as e:
            fused_chains = {"error": repr(e)}

    # ---- e2e: host buffers in, EVERY result back in host memory, copies inside the timed region.
    # Per step: x and y (2 x 64 MiB) up; the nine (n n) outputs (9 x 64 MiB) and the three reductions down.  Each case
    # writes its own device buffer; its download is queued on side branch 2 as soon as the kernel is issued, the next step's
    # upload runs on side branch 1 (PCIe is full duplex), and the step ends when both branches have joined.
    nb = n * n * 4
    big = [c for c, _ in sw.cases if not c.startswith("sum_")]
    hx = C.c_void_p(); hy = C.c_void_p(); hres = C.c_void_p(); hout = C.c_void_p()
    chk(lib.wf2cu_host_alloc(nb, C.byref(hx))); chk(lib.wf2cu_host_alloc(nb, C.byref(hy)))
    red_bytes = 2 * n * 4 + 4
    chk(lib.wf2cu_host_alloc(red_bytes, C.byref(hres)))
    chk(lib.wf2cu_host_alloc(len(big) * nb, C.byref(hout)))
    C.memmove(hx.value, sw.hx.ctypes.data, nb); C.memmove(hy.value, sw.hy.ctypes.data, nb)
    e2e_out = {c: wf.make_tensor((n, n)) for c in big}
    d2h_bytes = len(big) * nb + red_bytes

    def upload(s):
        chk(lib.wf2cu_h2d(s["x"].vec.ptr, 0, hx, nb))
        chk(lib.wf2cu_h2d(s["y"].vec.ptr, 0, hy, nb))

    def download(dev, host_off, nbytes):
        chk(lib.wf2cu_branch_begin(2))
        try:
            chk(lib.wf2cu_d2h_async(dev.vec.ptr, 0, hout.value + host_off if host_off >= 0 else hres.value - host_off - 1, nbytes))
        finally:
            chk(lib.wf2cu_branch_end(2))

    def e2e_run(steps):
        """`steps` sweeps; exactly `steps` uploads, `steps` sweeps and `steps` read-backs of every result."""
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
            for j, name in enumerate(big):
                s2 = dict(cur)
                s2["out"] = e2e_out[name]                      # (add_inplace accumulates into its own buffer from step to step)
                sw.run_case(name, s2)
                download(e2e_out[name], j * nb, nb)
            for name, key, off, nbytes in (("sum_axis1", "s1", 0, n * 4), ("sum_axis0", "s0", n * 4, n * 4), ("sum_all", "sa", 2 * n * 4, 4)):
                sw.run_case(name, cur)
                download(cur[key], -(off + 1), nbytes)
            chk(lib.wf2cu_branch_join(2))
            if i + 1 < steps:
                chk(lib.wf2cu_branch_join(1))
            sw.cursor = (sw.cursor + 1) % NSETS
            cur = nxt

    e2e_run(2)
    barrier()
    e2e_steps = max(3, min(args.steps, 10))
    chk(lib.wf2cu_timer_start())
    e2e_run(e2e_steps)
    t = C.c_float()
    chk(lib.wf2cu_timer_stop(C.byref(t)))
    e2e_ms = max_over_ranks(float(t.value)) / e2e_steps
    # what came back is what the device computed: the add result and the full sum, checked on the host copy
    host_add = np.ctypeslib.as_array(C.cast(hout.value, C.POINTER(C.c_float)), shape=(n * n,))[:4096].copy()
    host_sum = float(np.ctypeslib.as_array(C.cast(hres.value + 2 * n * 4, C.POINTER(C.c_float)), shape=(1,))[0])
    e2e_ok = bool(np.array_equal(host_add, (sw.hx + sw.hy).ravel()[:4096]) and
                  abs(host_sum - float(sw.hx.astype(np.float64).sum())) <= 1e-5 * float(np.abs(sw.hx).astype(np.float64).sum()))
    e2e = {"value": world * bytes_step / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e2e_ms,
           "h2d_bytes_per_step": 2 * nb, "d2h_bytes_per_step": d2h_bytes, "steps": e2e_steps, "results_checked_on_host": e2e_ok,
           "pcie_gbps_down": d2h_bytes / (e2e_ms * 1e-3) / 1e9,
           "api": "cl-waffe2_b200.ops (one libwaffe2cuda C-ABI call per node) + wf2cu_h2d / wf2cu_d2h_async with pinned host memory: every "
                  "result of the 12 cases is copied back (9 x 64 MiB + the three reductions); downloads queue on side branch 2 behind the "
                  "kernel that produced them, the upload of step i+1 runs on side branch 1 (full duplex); PCIe-bound"}
    for h in (hx, hy, hres, hout):
        chk(lib.wf2cu_host_free(h))
    e2e_out.clear()

    stage("e2e timed")
    # ---- MatmulNode sweep (C3), single GPU only
    gemm = None
    if args.gemm and world == 1:
        stage("gemm")
        gemm = gemm_sweep(wf, check=not args.no_cpu)
        if rank == 0 and not args.no_cpu:
            try:                                             # side information: may never break the bench line
                gemm["cpu_baseline"] = cpu_sgemm_baseline()
            except Exception as e:
                gemm["cpu_baseline"] = {"error": repr(e)}

    gpt2 = None
    if args.gpt2 and world == 1:
        stage("gpt2")
        try:
            gpt2 = gpt2_forward_bench(wf, oracle_check=not args.no_cpu)
        except Exception as e:                       # reported, never silently dropped
            gpt2 = {"error": repr(e)}

    cnn = None
    if args.cnn and world == 1:
        stage("cnn")
        try:
            cnn = importlib.import_module("cl-waffe2_b200.mlp_bench").run_cnn(wf)
        except Exception as e:
            cnn = {"error": repr(e)}

    # ---- MLP training step (the sharded path)
    mlp = mlp_detail = None
    if args.mlp:
        stage("mlp")
        mlp_mod = importlib.import_module("cl-waffe2_b200.mlp_bench")
        mlp, mlp_detail = mlp_mod.run(wf, dist, rank, world, steps=args.mlp_steps, warmup=args.mlp_warmup, exchange=args.exchange,
                                      host_losses=None if args.no_cpu else mlp_host_losses,
                                      cpu_fn=None if args.no_cpu else mlp_cpu_baseline)
        for k in ("weak", "strong", "weak_nccl", "n1_same_run"):
            if mlp_detail.get(k):
                mlp_detail[k].pop("loss_first20", None)
    stage("mlp done")

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

    # a library GEMM as the second tensor denominator: after every number of this repo's own kernels has been taken
    if gemm is not None and args.cublas:
        stage("cublas tf32")
        try:
            gemm["cublas_tf32"] = cublas_tf32_peak()
            top = gemm["roofline"]
            gemm["roofline"]["pipe_frac_of_cublas_tf32"] = top["achieved"] * top["pipe_units"] / gemm["cublas_tf32"]["tflops"]
            gemm["roofline"]["fp32_equivalent_vs_cublas_tf32"] = top["achieved"] / gemm["cublas_tf32"]["tflops"]
        except Exception as e:
            gemm["cublas_tf32"] = {"error": repr(e)}

    library = None
    if world == 1 and args.cublas:
        stage("torch library cases")
        try:
            library = torch_library_cases(n)
            for k, v in library.items():
                if isinstance(v, dict) and k in per_case:
                    v["ours_ms"] = per_case[k]["ms"]
                    v["ours_speedup"] = v["ms"] / per_case[k]["ms"]
        except Exception as e:
            library = {"error": repr(e)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
            "warmup_requested": args.warmup,
            "ms_per_step": ms_per_step, "eager_ms_per_step": eager_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(n, sw.cases),
            "method": {"launch": "one step captured into a CUDA graph and replayed" + (
                           f"; the 12 cases are independent nodes (each has its own buffer set) and alternate between {nbr} side branches of the "
                           "graph (wf2cu_branch_*), so the launch boundary of one chain overlaps the streaming of the other; "
                           "single_chain = the same step with every case serialised" if nbr > 1 else ""),
                       "branches": nbr,
                       "single_chain": {"ms_per_step": single_ms, "value": world * bytes_step / (single_ms * 1e-3) / 1e9},
                       "parallelism": "replicas only (the sharded path is the `mlp` object)" if world > 1 else "single GPU"},
            "roofline": roofline, "cases": per_case, "library_cases": library, "fused_chains": fused_chains, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": int(launches1.value - launches0.value), "clocks": clocks.summary(),
            "gemm": gemm, "gpt2": gpt2, "cnn": cnn, "mlp_detail": mlp_detail,
            "mlp": mlp,                                       # compact and LAST: survives the driver's tail at every N
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
    ap.add_argument("--fused", type=int, default=1, help="fused_chains block (compiled element-wise chains + Adam arena)")
    ap.add_argument("--branches", type=int, default=2, help="independent cases of the sweep step alternate between this many side branches (1 = one chain)")
    ap.add_argument("--cublas", type=int, default=1, help="measure cuBLAS TF32 at 8192^3 through torch as a second GEMM denominator")
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
