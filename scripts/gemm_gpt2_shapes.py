#!/usr/bin/env python
"""The four linear layers of a GPT-2 small block at B = 8, N = 1024 (8192 rows) through wf2cu_gemm_ex2 with the weight
declared constant (as gpt2.py does): time per call for the three persistent tile shapes and auto, all launches included."""
import ctypes as C, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
rng = np.random.default_rng(0)
modes = [int(x) for x in os.environ.get("MODES", "2,5,6,0").split(",")]
reps = int(os.environ.get("REPS", 4))
for name, (m, n, k) in {"qkv": (8192, 2304, 768), "attn_proj": (8192, 768, 768), "fc": (8192, 3072, 768), "fc_proj": (8192, 768, 3072)}.items():
    A = wf.from_numpy(rng.uniform(-1, 1, size=(m, k)).astype(np.float32))
    W = wf.from_numpy(rng.uniform(-1, 1, size=(k, n)).astype(np.float32))
    b = wf.from_numpy(rng.uniform(-1, 1, size=(n,)).astype(np.float32))
    o = wf.make_tensor((m, n))
    chk(lib.wf2cu_gemm_const_operand(W.vec.ptr, 0, k * n, 1))
    line = f"{name:9s} m={m} n={n} k={k}:"
    for mode in modes:
        chk(lib.wf2cu_gemm_set_mode(mode))
        f = lambda: chk(lib.wf2cu_gemm_ex2(0, 0, m, n, k, A.vec.ptr, 0, k, W.vec.ptr, 0, n, o.vec.ptr, 0, n, b.vec.ptr, 0, 0, None, 0, 0))
        f(); f(); wf.ops.sync()
        best = 1e9
        for _ in range(5):
            chk(lib.wf2cu_timer_start())
            for _ in range(reps):
                f()
            ms = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(ms)))
            best = min(best, ms.value / reps)
        path = C.c_int(0); chk(lib.wf2cu_gemm_last_path(C.byref(path)))
        line += f"  mode{mode}: {best * 1e3:6.1f} us {2.0 * m * n * k / best / 1e9:5.1f} TF (path {path.value})"
    print(line, flush=True)
    chk(lib.wf2cu_gemm_const_operand(W.vec.ptr, 0, k * n, 0))
chk(lib.wf2cu_gemm_set_mode(0))
