#!/usr/bin/env python
"""GPU-box script: the MLP-step GEMM shapes (C4) through wf2cu_gemm in mode 1 (fp32 SIMT, incl. the cluster
split-K kernel) and mode 2 (tcgen05 3xTF32) — parity vs fp64 and time per call inside a CUDA graph."""
import ctypes as C, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
rng = np.random.default_rng(0)
# (m, n, k, ta, tb) as the MLP step issues them: x @ W^T, dout @ W, dout^T @ x
shapes = [(1024, 256, 784, 0, 1), (1024, 256, 256, 0, 1), (1024, 10, 256, 0, 1),
          (1024, 256, 10, 0, 0), (10, 256, 1024, 1, 0), (1024, 256, 256, 0, 0), (256, 256, 1024, 1, 0), (256, 784, 1024, 1, 0),
          (128, 256, 784, 0, 1), (128, 784, 256, 0, 0), (256, 784, 128, 1, 0),
          (100, 10, 100, 0, 1), (100, 1, 10, 0, 1), (3, 5, 4, 0, 0), (130, 70, 33, 1, 1), (64, 64, 4096, 0, 0), (200, 300, 555, 1, 0),
          (512, 512, 512, 0, 0), (1024, 1024, 1024, 0, 0), (7, 3, 5, 1, 1), (129, 65, 31, 1, 0), (127, 63, 33, 0, 1), (300, 200, 1000, 1, 1)]
MODES = (1, 4, 0)
if os.environ.get("WF2_SHAPES") == "mid":
    # where does the one-launch fused tcgen05 kernel (mode 4) stop beating planes + persistent kernel (mode 2)?
    shapes = [(768, 768, 768, 0, 0), (1024, 1024, 1024, 0, 0), (1280, 1280, 1280, 0, 0), (1536, 1536, 1536, 0, 0),
              (2048, 2048, 2048, 0, 0), (3072, 3072, 3072, 0, 0), (8192, 2304, 768, 0, 0), (8192, 768, 768, 0, 0),
              (8192, 3072, 768, 0, 0), (8192, 768, 3072, 0, 0), (1024, 1024, 64, 0, 0), (1024, 64, 1024, 0, 0)]
    MODES = (2, 4, 0)
bad = 0
for (m, n, k, ta, tb) in shapes:
    A = rng.uniform(-1, 1, size=((k, m) if ta else (m, k))).astype(np.float32)
    B = rng.uniform(-1, 1, size=((n, k) if tb else (k, n))).astype(np.float32)
    truth = (A.T if ta else A).astype(np.float64) @ (B.T if tb else B).astype(np.float64)
    ga, gb = wf.from_numpy(A), wf.from_numpy(B)
    o = wf.make_tensor((m, n))
    line = f"m={m:5d} n={n:5d} k={k:5d} ta={ta} tb={tb}:"
    for mode in MODES:
        chk(lib.wf2cu_gemm_set_mode(mode))
        try:
            f = lambda: wf.ops.matmul(wf.ops.t(ga) if ta else ga, wf.ops.t(gb) if tb else gb, out=o)
            f(); f()
        except Exception as e:
            line += f"  mode{mode}: declined"
            continue
        err = float(np.abs(o.to_numpy() - truth).max() / np.abs(truth).max())
        ok = err <= 2e-5
        bad += 0 if ok else 1
        ex = C.c_void_p()
        chk(lib.wf2cu_graph_begin())
        for _ in range(20): f()
        chk(lib.wf2cu_graph_end(C.byref(ex)))
        chk(lib.wf2cu_graph_launch(ex))
        best = 1e9
        for _ in range(3):
            chk(lib.wf2cu_timer_start()); chk(lib.wf2cu_graph_launch(ex))
            ms = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(ms)))
            best = min(best, ms.value / 20)
        chk(lib.wf2cu_graph_destroy(ex))
        line += f"  mode{mode}: {best*1e3:7.2f} us err {err:.1e}{'' if ok else ' FAIL'}"
    print(line, flush=True)
chk(lib.wf2cu_gemm_set_mode(0))
print("failures:", bad)
sys.exit(1 if bad else 0)
