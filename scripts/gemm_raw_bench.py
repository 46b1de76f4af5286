#!/usr/bin/env python
"""MatmulNode: the raw-operand kernel (mode 7, no operand pre-pass) against the planes kernels (modes 2 / 5 / 6) and auto,
every launch of a call timed, on the C3 squares and on GPT-2's projection shapes."""
import ctypes as C, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
lib = wf._lib.load(); chk = wf._lib.check
rng = np.random.default_rng(0)


def timed(fn, reps):
    fn(); fn(); chk(lib.wf2cu_sync())
    best = None
    for _ in range(3):
        chk(lib.wf2cu_timer_start())
        for _ in range(reps):
            fn()
        ms = C.c_float(); chk(lib.wf2cu_timer_stop(C.byref(ms)))
        best = ms.value / reps if best is None else min(best, ms.value / reps)
    return best * 1e3


shapes = [(2048, 2048, 2048, False), (4096, 4096, 4096, False), (4096, 4096, 4096, True), (8192, 8192, 8192, False),
          (8192, 2304, 768, False), (8192, 768, 768, False), (8192, 3072, 768, False), (8192, 768, 3072, False), (1024, 256, 784, True)]
print(f"{'m':>6} {'n':>6} {'k':>6} {'B':>3} | " + " ".join(f"{'mode ' + str(md):>10}" for md in (0, 7, 5, 6, 2)) + "   (us per MatmulNode, all launches)")
for m, n, k, tb in shapes:
    A = wf.from_numpy(rng.uniform(-1, 1, size=(m, k)).astype(np.float32))
    B = wf.from_numpy(rng.uniform(-1, 1, size=((n, k) if tb else (k, n))).astype(np.float32))
    o = wf.make_tensor((m, n))
    row = []
    for md in (0, 7, 5, 6, 2):
        chk(lib.wf2cu_gemm_set_mode(md))
        try:
            t = timed(lambda: wf.ops.matmul(A, wf.ops.t(B) if tb else B, out=o), 10 if m * n * k < 1e11 else 4)
            row.append(f"{t:10.1f}")
        except wf.Wf2CudaError:
            row.append(f"{'-':>10}")
        finally:
            chk(lib.wf2cu_gemm_set_mode(0))
    print(f"{m:6d} {n:6d} {k:6d} {'NT' if tb else 'NN':>3} | " + " ".join(row) + f"   raw: {2.0 * m * n * k / float(row[1]) / 1e6:.0f} TFLOP/s" if row[1].strip() != '-' else "")
    del A, B, o
