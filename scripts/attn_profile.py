#!/usr/bin/env python
"""The two products of GPT-2's attention at B = 8, N = 1024 (96 heads): Q.K^T and P.V — the program ncu wraps."""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
rng = np.random.default_rng(0)
Q = wf.from_numpy(rng.uniform(-1, 1, size=(96, 1024, 64)).astype(np.float32))
K = wf.from_numpy(rng.uniform(-1, 1, size=(96, 1024, 64)).astype(np.float32))
V = wf.from_numpy(rng.uniform(-1, 1, size=(96, 1024, 64)).astype(np.float32))
w = wf.make_tensor((96, 1024, 1024))
o = wf.make_tensor((96, 1024, 64))
def timed(f, n=4):
    f(); f(); wf.ops.sync()
    best = 1e9
    for _ in range(5):
        chk(lib.wf2cu_timer_start())
        for _ in range(n): f()
        ms = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(ms)))
        best = min(best, ms.value / n)
    return best * 1e3
t1 = timed(lambda: wf.ops.matmul(Q, wf.ops.t(K), out=w))
t2 = timed(lambda: wf.ops.matmul(w, V, out=o))
print(f"QK^T {t1:.1f} us   P.V {t2:.1f} us ({96*1024*1024*4/t2/1e6:.0f} GB/s of P)")
