#!/usr/bin/env python
"""The GPT-2 example's attention block at B = 8, N = 1024, 12 heads of 64 (examples/gpt-2/utils.lisp:222-244) as the two
launches the library provides — wf2cu_attn_heads_softmax (scores / N, softmax over heads, P written once) and wf2cu_attn_pv
(P.V, merged heads) — the program ncu wraps."""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
rng = np.random.default_rng(0)
B, N, H, D = 8, 1024, 12, 64
qkv = wf.from_numpy(rng.normal(0, 1, size=(B, N, 3 * H * D)).astype(np.float32))
P = wf.make_tensor((B, H, N, N))
out = wf.make_tensor((B, N, H * D))
def scores(): chk(lib.wf2cu_attn_heads_softmax(qkv.vec.ptr, 0, B, N, H, D, float(N), P.vec.ptr, 0))
def pv(): chk(lib.wf2cu_attn_pv(P.vec.ptr, 0, qkv.vec.ptr, 0, B, N, H, D, out.vec.ptr, 0))
def timed(f, n=4):
    f(); f(); wf.ops.sync()
    best = 1e9
    for _ in range(5):
        chk(lib.wf2cu_timer_start())
        for _ in range(n): f()
        ms = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(ms)))
        best = min(best, ms.value / n)
    return best * 1e3
t1, t2 = timed(scores), timed(pv)
pbytes = B * H * N * N * 4
print(f"heads-softmax {t1:.1f} us ({pbytes / t1 / 1e6:.0f} GB/s of P written)   P.V {t2:.1f} us ({pbytes / t2 / 1e6:.0f} GB/s of P read)")
