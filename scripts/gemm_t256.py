#!/usr/bin/env python
"""A/B of the persistent tcgen05 kernels: mode 2 (128x256 tiles), 5 (256x256), 6 (256x192), parity vs fp64 + time."""
import ctypes as C, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
rng = np.random.default_rng(0)
shapes = [(512, 512, 256, 0, 0), (2048, 2048, 2048, 0, 0), (4096, 4096, 4096, 0, 0), (4096, 4096, 4096, 1, 1), (8192, 8192, 8192, 0, 0),
          (8192, 2304, 768, 0, 0), (8192, 3072, 768, 0, 0), (8192, 768, 3072, 0, 0), (8192, 768, 768, 0, 0), (8192, 1536, 1536, 0, 0), (3000, 2500, 1001, 0, 1), (300, 700, 100, 1, 0)]
if len(sys.argv) > 1 and sys.argv[1] == "quick":
    shapes = shapes[:1]
for (m, n, k, ta, tb) in shapes:
    A = rng.uniform(-1, 1, size=((k, m) if ta else (m, k))).astype(np.float32)
    B = rng.uniform(-1, 1, size=((n, k) if tb else (k, n))).astype(np.float32)
    truth = None
    if m * n * k <= 4096 ** 3:
        truth = (A.T if ta else A).astype(np.float64) @ (B.T if tb else B).astype(np.float64)
    ga, gb = wf.from_numpy(A), wf.from_numpy(B)
    o = wf.make_tensor((m, n))
    line = f"m={m:5d} n={n:5d} k={k:5d} ta={ta} tb={tb}:"
    outs = {}
    for mode in (2, 5, 6, 0):
        chk(lib.wf2cu_gemm_set_mode(mode))
        f = lambda: wf.ops.matmul(wf.ops.t(ga) if ta else ga, wf.ops.t(gb) if tb else gb, out=o)
        f(); f()
        wf.ops.sync()
        outs[mode] = o.to_numpy()
        err = float(np.abs(outs[mode] - truth).max() / np.abs(truth).max()) if truth is not None else float("nan")
        best = 1e9
        for _ in range(5):
            chk(lib.wf2cu_timer_start()); f(); f(); f(); f()
            ms = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(ms)))
            best = min(best, ms.value / 4)
        tf = 2.0 * m * n * k / (best * 1e-3) / 1e12
        line += f"  m{mode}: {best*1e3:7.1f} us {tf:5.1f} TF {err:.0e}"
    line += f"  same={np.array_equal(outs[2], outs[5]) and np.array_equal(outs[2], outs[6])}"
    print(line, flush=True)
chk(lib.wf2cu_gemm_set_mode(0))
# attention's Q.K^T: 96 heads of (1024 64).(64 1024), K operand stored (1024 64) -> transposed flag
Q = wf.from_numpy(rng.uniform(-1, 1, size=(96, 1024, 64)).astype(np.float32))
Kt = wf.from_numpy(rng.uniform(-1, 1, size=(96, 1024, 64)).astype(np.float32))
o = wf.make_tensor((96, 1024, 1024))
for mode in (2, 5, 0):
    chk(lib.wf2cu_gemm_set_mode(mode))
    f = lambda: wf.ops.matmul(Q, wf.ops.t(Kt), out=o)
    f(); f(); wf.ops.sync()
    best = 1e9
    for _ in range(5):
        chk(lib.wf2cu_timer_start()); f(); f(); f(); f()
        ms = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(ms)))
        best = min(best, ms.value / 4)
    print(f"QK^T 96x(1024 64)(64 1024) mode{mode}: {best*1e3:8.1f} us  ({96*1024*1024*4/best/1e6:.0f} GB/s of output)", flush=True)
chk(lib.wf2cu_gemm_set_mode(0))
