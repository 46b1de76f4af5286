#!/usr/bin/env python
"""Time the GPT-2 example's attention block at the benchmarked shape (B = 8, N = 1024, 12 heads x 64): the two fused kernels
(csrc/attention.cu) against the node sequence they replace, each under a CUDA graph of 5 calls."""
import ctypes as C, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
wf = importlib.import_module("cl-waffe2_b200")
g = importlib.import_module("cl-waffe2_b200.gpt2")
wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
B, N, H, D = int(os.environ.get("B", 8)), int(os.environ.get("N", 1024)), 12, 64
rng = np.random.default_rng(0)
x = wf.from_numpy((rng.normal(size=(B, N, 3 * H * D)) * 2).astype(np.float32))
P = wf.make_tensor([B, H, N, N]); out = wf.make_tensor([B, N, H * D])


def timed(fn, reps=5):
    fn(); chk(lib.wf2cu_sync())
    best = None
    for _ in range(3):
        chk(lib.wf2cu_timer_start())
        for _ in range(reps):
            fn()
        ms = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(ms)))
        best = ms.value / reps if best is None else min(best, ms.value / reps)
    return best * 1e3


def scores(): chk(lib.wf2cu_attn_heads_softmax(x.vec.ptr, 0, B, N, H, D, float(N), P.vec.ptr, 0))
def pv(): chk(lib.wf2cu_attn_pv(P.vec.ptr, 0, x.vec.ptr, 0, B, N, H, D, out.vec.ptr, 0))


pool = wf.StaticPool()
def nodes():
    with pool:
        g.self_attention(x, H)
g.FUSE_ATTENTION = False
nodes(); nodes(); pool.frozen = True
t_nodes = timed(nodes)
g.FUSE_ATTENTION = True
t_s, t_pv = timed(scores), timed(pv)
pbytes = B * H * N * N * 4
print(f"attention B={B} N={N}: heads-softmax {t_s:.1f} us ({pbytes / t_s / 1e3:.0f} GB/s of P written), P.V {t_pv:.1f} us "
      f"({pbytes / t_pv / 1e3:.0f} GB/s of P read), fused block {t_s + t_pv:.1f} us; node sequence {t_nodes:.1f} us")
