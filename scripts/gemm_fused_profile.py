#!/usr/bin/env python
"""Two MLP-step GEMMs through the fused tcgen05 kernel (mode 4) — the program ncu wraps."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
rng = np.random.default_rng(0)
chk(lib.wf2cu_gemm_set_mode(4))
for (m, n, k, ta, tb) in [(1024, 256, 784, 0, 1), (10, 256, 1024, 1, 0), (256, 784, 1024, 1, 0)]:
    A = wf.from_numpy(rng.uniform(-1, 1, size=((k, m) if ta else (m, k))).astype(np.float32))
    B = wf.from_numpy(rng.uniform(-1, 1, size=((n, k) if tb else (k, n))).astype(np.float32))
    o = wf.make_tensor((m, n))
    for _ in range(2):
        wf.ops.matmul(wf.ops.t(A) if ta else A, wf.ops.t(B) if tb else B, out=o)
wf.ops.sync()
print("ok")
