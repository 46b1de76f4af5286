#!/usr/bin/env python
"""Generate tests/golden/ref_kernels_v1.npz from the REFERENCE's own leaf kernels.

Runs in the authoring container only: it needs oracle/_ref/libwaffe2simd_ref.so, i.e. cl-waffe2's
`cl-waffe2-simd/kernels/cl-waffe2-simd.c` + `im2col.c` + the vendored SLEEF compiled from where they lie under
/root/reference by oracle/Makefile.  The fixture travels with the repository, so the parity tests stay anchored on the
reference's outputs on machines that have neither /root/reference nor the compiled reference library.

    python tests/golden/make_golden.py          # rewrites ref_kernels_v1.npz (deterministic: fixed seeds)

Contents (all float32 unless noted):
  unary/<fn>/x, unary/<fn>/y          waffe2_s<fn>(n, x, 1, y, 1), 520 points over the domain the GPU parity test uses
  arith/x, arith/y, arith/<op>        waffe2_s{add,sub,mul,div}: x <- x op y
  pow/x, pow/<n>                      waffe2_spow for n in 2, 3, 0.5, -1.5
  cmp/x, cmp/y, cmp/<op>, cmp/<op>_scal   waffe2_s<op>(x, y -> out, then 2, else -1); <op>_scal against 0.25, then 1 else 0
  maxmin/x, maxmin/max, maxmin/min    waffe2_s{max,min} over 520 elements
  seqsum/x, seqsum/sum                `!sum`'s leaf call: waffe2_sadd(n, acc, incx=0, x, 1): the sequential fp32 sum
  bcast/x, bcast/y, bcast/mul         waffe2_smul(n, x, 1, y, incy=0): a (:broadcast n) operand
  unfold/img, unfold/col, unfold/back waffe2_im2col_s_column / waffe2_col2im_s_column on (2 3 7 8), 3x3, stride 2 1, pad 0
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import kernels  # noqa: E402

OUT = os.path.join(HERE, "ref_kernels_v1.npz")
N = 520                      # 65 AVX2 vectors: the SIMD body; the scalar tail is exercised by N_TAIL
N_TAIL = 5
UNARY_DOMAINS = {
    "sin": (-20, 20), "cos": (-20, 20), "tan": (-1.5, 1.5), "asin": (-1, 1), "acos": (-1, 1), "atan": (-50, 50),
    "sinh": (-8, 8), "cosh": (-8, 8), "tanh": (-6, 6), "exp": (-30, 30), "log": (1e-6, 1e4),
    "log2": (1e-6, 1e4), "log10": (1e-6, 1e4), "log1p": (-0.9, 100), "sqrt": (0, 1e6), "abs": (-5, 5),
}


def main():
    if not kernels.have_reference():
        raise SystemExit("oracle/_ref/libwaffe2simd_ref.so is missing: run `make -C oracle all` where /root/reference exists")
    K = kernels.load("reference")
    assert K.kind == "reference"
    rng = np.random.default_rng(20260)
    out = {}
    n = N + N_TAIL
    for fn, (lo, hi) in sorted(UNARY_DOMAINS.items()):
        x = rng.uniform(lo, hi, n).astype(np.float32)
        y = np.zeros(n, np.float32)
        K.unary(fn, n, x, 0, 1, y, 0, 1)
        out[f"unary/{fn}/x"], out[f"unary/{fn}/y"] = x, y
    x = rng.uniform(-4, 4, n).astype(np.float32)
    y = rng.uniform(0.25, 4, n).astype(np.float32) * rng.choice(np.array([-1, 1], np.float32), n)
    out["arith/x"], out["arith/y"] = x, y
    for op in kernels.ARITH:
        acc = x.copy()
        K.arith(op, n, acc, 0, 1, y, 0, 1)
        out[f"arith/{op}"] = acc
    px = rng.uniform(0.1, 4, n).astype(np.float32)
    out["pow/x"] = px
    for p in (2.0, 3.0, 0.5, -1.5):
        o = np.zeros(n, np.float32)
        K.pow(n, px, 0, 1, p, o, 0, 1)
        out[f"pow/{p}"] = o
    cx = rng.integers(-2, 3, n).astype(np.float32) * np.float32(0.25)
    cy = rng.integers(-2, 3, n).astype(np.float32) * np.float32(0.25)
    out["cmp/x"], out["cmp/y"] = cx, cy
    for op in kernels.CMP:
        o = np.zeros(n, np.float32)
        K.cmp(op, n, cx, 0, 1, cy, 0, 1, o, 0, 1, 2.0, -1.0)
        out[f"cmp/{op}"] = o
        o = np.zeros(n, np.float32)
        K.cmp_scal(op, n, cx, 0, 1, o, 0, 1, 0.25, 1.0, 0.0)
        out[f"cmp/{op}_scal"] = o
    mx = rng.normal(size=n).astype(np.float32)
    out["maxmin/x"] = mx
    for name in ("max", "min"):
        o = np.zeros(1, np.float32)
        K.maxmin(name, n, mx, 0, 1, o, 0)
        out[f"maxmin/{name}"] = o
    sx = rng.uniform(-1, 1, 4099).astype(np.float32)
    acc = np.zeros(1, np.float32)
    K.arith("add", len(sx), acc, 0, 0, sx, 0, 1)            # out*[:broadcast] += x  (reduction.lisp:60-79)
    out["seqsum/x"], out["seqsum/sum"] = sx, acc
    bx = rng.uniform(-2, 2, n).astype(np.float32)
    by = np.array([1.7], np.float32)
    acc = bx.copy()
    K.arith("mul", n, acc, 0, 1, by, 0, 0)
    out["bcast/x"], out["bcast/y"], out["bcast/mul"] = bx, by, acc
    # im2col / col2im (im2col.c:63-115, 395-448): (N C H W) = (2 3 7 8), kernel 3x3, stride (2 1), no padding, dilation 1
    Nn, C, H, W, KH, KW, sh, sw = 2, 3, 7, 8, 3, 3, 2, 1
    Ho, Wo = (H - KH) // sh + 1, (W - KW) // sw + 1
    img = rng.normal(size=Nn * C * H * W).astype(np.float32)
    col = np.zeros(Nn * C * KH * KW * Ho * Wo, np.float32)
    K.unfold("im2col", col, Nn, C, H, W, Ho, Wo, KH, KW, 0, 0, sh, sw, 1, 1, img)
    back = np.zeros_like(img)
    K.unfold("col2im", col, Nn, C, H, W, Ho, Wo, KH, KW, 0, 0, sh, sw, 1, 1, back)
    out["unfold/img"], out["unfold/col"], out["unfold/back"] = img, col, back
    out["unfold/geometry"] = np.array([Nn, C, H, W, Ho, Wo, KH, KW, 0, 0, sh, sw, 1, 1], np.int64)
    np.savez_compressed(OUT, **out)
    print(f"wrote {OUT}: {len(out)} arrays, {os.path.getsize(OUT)} bytes")


if __name__ == "__main__":
    main()
