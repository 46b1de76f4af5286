#!/usr/bin/env python
"""A/B of the two persistent tcgen05 kernels: mode 2 (128x256 tiles) vs mode 5 (256x256 tiles), parity vs fp64 + time."""
import ctypes as C, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
rng = np.random.default_rng(0)
shapes = [(512, 512, 256, 0, 0), (2048, 2048, 2048, 0, 0), (4096, 4096, 4096, 0, 0), (4096, 4096, 4096, 1, 1), (8192, 8192, 8192, 0, 0),
          (8192, 2304, 768, 0, 0), (8192, 3072, 768, 0, 0), (8192, 768, 3072, 0, 0), (3000, 2500, 1001, 0, 1), (300, 700, 100, 1, 0)]
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
    for mode in (2, 5):
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
        line += f"  mode{mode}: {best*1e3:8.1f} us {tf:6.1f} TF err {err:.1e}"
    line += f"  same={np.array_equal(outs[2], outs[5])} maxdiff={float(np.abs(outs[2]-outs[5]).max()):.2e}"
    print(line, flush=True)
chk(lib.wf2cu_gemm_set_mode(0))
