#!/usr/bin/env python
"""CPU emulation of the two split products MatmulNode's plane-fed kernels can run (csrc/gemm_tc.cu `split_bf`, `split1`), with
fp64 accumulation so that only the SPLIT's error shows (the kernels add the fp32 accumulation in tensor memory on top):

    3xTF32          C = hi.hi + hi.tf32(lo) + tf32(lo).hi           hi = tf32_rn(x), lo = x - hi
    TF32 + 2 x bf16 C = hi.hi + bf16(x).bf16(lo) + bf16(lo).bf16(x)

against an fp64 product, relative to max|C|, beside numpy's own fp32 matmul.  Prints one line per distribution and size."""
import numpy as np
import torch


def tf32_rn(x):
    u = x.view(np.uint32).astype(np.uint64)
    u = (u + 0xFFF + ((u >> 13) & 1)) & 0xFFFFE000
    return u.astype(np.uint32).view(np.float32)


def bf16_rn(x):
    return torch.from_numpy(np.ascontiguousarray(x)).to(torch.bfloat16).to(torch.float32).numpy()


def products(A, B):
    ah, bh = tf32_rn(A), tf32_rn(B)
    al, bl = A - ah, B - bh
    d = np.float64
    main = ah.astype(d) @ bh.astype(d)
    c3 = main + ah.astype(d) @ tf32_rn(bl).astype(d) + tf32_rn(al).astype(d) @ bh.astype(d)
    ch = main + bf16_rn(A).astype(d) @ bf16_rn(bl).astype(d) + bf16_rn(al).astype(d) @ bf16_rn(B).astype(d)
    return c3, ch


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    print(f"{'operands':>10} {'k':>5} | {'3xTF32':>9} {'TF32+2bf16':>11} {'numpy fp32':>11}   (max |error| / max|C|)")
    for dist in ("uniform", "normal", "positive", "integers"):
        for k in (1024, 4096):
            m = 256
            if dist == "uniform":
                A, B = rng.uniform(-1, 1, (m, k)), rng.uniform(-1, 1, (k, m))
            elif dist == "normal":
                A, B = rng.normal(0, 1, (m, k)), rng.normal(0, 1, (k, m))
            elif dist == "positive":
                A, B = rng.uniform(0, 1, (m, k)), rng.uniform(0, 1, (k, m))
            else:
                A, B = rng.integers(-8, 9, (m, k)), rng.integers(-8, 9, (k, m))
            A, B = A.astype(np.float32), B.astype(np.float32)
            truth = A.astype(np.float64) @ B.astype(np.float64)
            c3, ch = products(A, B)
            den = np.abs(truth).max()
            print(f"{dist:>10} {k:5d} | {np.abs(c3 - truth).max() / den:9.2e} {np.abs(ch - truth).max() / den:11.2e} "
                  f"{np.abs((A @ B) - truth).max() / den:11.2e}")
