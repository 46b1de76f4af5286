#!/usr/bin/env python
"""The CTA-pair kernel (mode 8: tcgen05.mma.cta_group::2, 512x256 pair tiles, half of B per CTA) against the single-CTA plane
kernels (modes 5 / 6 / 2) and auto: time per MatmulNode (every launch of the call), error against an fp64 product — over ALL
rows for the first shape (so a swapped half or a missed tile cannot hide), 64 sampled rows otherwise."""
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


shapes = [(1000, 517, 333, False), (1024, 1024, 1024, False), (2048, 2048, 2048, False), (4096, 4096, 4096, False), (4096, 4096, 4096, True),
          (8192, 8192, 8192, False), (8192, 2304, 768, False), (8192, 768, 768, False), (8192, 3072, 768, False), (8192, 768, 3072, False)]
if len(sys.argv) > 1:
    shapes = shapes[:int(sys.argv[1])]
modes = (8, 5, 6, 2, 0)
print(f"{'m':>6} {'n':>6} {'k':>6} {'B':>3} {'corr':>5} | " + " ".join(f"{'mode ' + str(md):>17}" for md in modes) + "   (us per MatmulNode ; err / max|C| ; path)", flush=True)
for si, (m, n, k, tb) in enumerate(shapes):
    a = rng.uniform(-1, 1, size=(m, k)).astype(np.float32)
    b = rng.uniform(-1, 1, size=((n, k) if tb else (k, n))).astype(np.float32)
    rows = np.arange(m) if si < 2 else rng.choice(m, size=64, replace=False)
    truth = a[rows].astype(np.float64) @ (b.T if tb else b).astype(np.float64)
    den = np.abs(truth).max()
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
            except wf.Wf2CudaError as e:
                row.append(f"{'-':>17}")
                if md == 8: print("   mode 8:", str(e)[:200], flush=True)
            finally:
                chk(lib.wf2cu_gemm_set_mode(0))
        print(f"{m:6d} {n:6d} {k:6d} {'NT' if tb else 'NN':>3} {name:>5} | " + " ".join(row) +
              f"   pair {2.0 * m * n * k / float(row[0].split()[0]) / 1e6 if row[0].strip() != '-' else 0:.0f} TFLOP/s", flush=True)
    chk(lib.wf2cu_gemm_set_correction(0))
    del A, B, o
