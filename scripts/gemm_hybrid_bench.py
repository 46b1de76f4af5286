#!/usr/bin/env python
"""MatmulNode on the plane-fed tcgen05 kernels: TF32 + two bf16 correction MMAs (the default) against the classic 3xTF32 form
(wf2cu_gemm_set_correction), per tile shape (modes 2 / 5 / 6 = 128x256 / 256x256 / 256x192) and auto, every launch of a call
timed; error of each against an fp64 product on 64 sampled rows, relative to max|C|, with numpy's fp32 matmul beside it."""
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
          (8192, 2304, 768, False), (8192, 768, 768, False), (8192, 3072, 768, False), (8192, 768, 3072, False)]
if len(sys.argv) > 1:
    shapes = shapes[:int(sys.argv[1])]
modes = (0, 5, 6, 2)
print(f"{'m':>6} {'n':>6} {'k':>6} {'B':>3} {'corr':>5} | " + " ".join(f"{'mode ' + str(md):>17}" for md in modes) + "   (us per MatmulNode ; err / max|C|)")
for m, n, k, tb in shapes:
    a = rng.uniform(-1, 1, size=(m, k)).astype(np.float32)
    b = rng.uniform(-1, 1, size=((n, k) if tb else (k, n))).astype(np.float32)
    rows = rng.choice(m, size=64, replace=False)
    truth = a[rows].astype(np.float64) @ (b.T if tb else b).astype(np.float64)
    den = np.abs(truth).max()
    f32 = np.abs((a[rows] @ (b.T if tb else b)) - truth).max() / den
    A, B, o = wf.from_numpy(a), wf.from_numpy(b), wf.make_tensor((m, n))
    for corr, name in ((0, "bf16"), (1, "tf32")):
        chk(lib.wf2cu_gemm_set_correction(corr))
        row = []
        for md in modes:
            chk(lib.wf2cu_gemm_set_mode(md))
            try:
                fn = lambda: wf.ops.matmul(A, wf.ops.t(B) if tb else B, out=o)
                t = timed(fn, 10 if m * n * k < 1e11 else 4)
                path = C.c_int(); chk(lib.wf2cu_gemm_last_path(C.byref(path)))
                err = np.abs(o.to_numpy()[rows] - truth).max() / den
                row.append(f"{t:8.1f} {err:7.1e}p{path.value}")
            except wf.Wf2CudaError:
                row.append(f"{'-':>17}")
            finally:
                chk(lib.wf2cu_gemm_set_mode(0))
        best = min(float(r.split()[0]) for r in row if r.strip() != '-')
        print(f"{m:6d} {n:6d} {k:6d} {'NT' if tb else 'NN':>3} {name:>5} | " + " ".join(row) +
              f"   best {2.0 * m * n * k / best / 1e6:.0f} TFLOP/s  (numpy fp32 err {f32:.1e})")
    chk(lib.wf2cu_gemm_set_correction(0))
    del A, B, o
