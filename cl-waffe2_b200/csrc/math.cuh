// math.cuh — device math for the unary nodes (`source/base-impl/mathematics.lisp:13-257`).
// The reference evaluates these with SLEEF u10 (<= 1.0 ulp) AVX2 entry points
// (`cl-waffe2-simd/kernels/simd/x86_64/avx2.h:264-332`), and with double-precision libm
// narrowed to float on the n%8 tail (`cl-waffe2-simd.c:636-641`).  Parity bar: <= 2 ulp against
// either.  Functions whose CUDA fp32 implementation is documented at <= 1 ulp are used as is
// (compiled WITHOUT --use_fast_math); the rest are evaluated in fp64 and narrowed, which is the
// reference's own tail arithmetic.
#pragma once
#include <cuda_runtime.h>
#include "../../include/waffe2cuda.h"

namespace wf2 {

template <int FN> struct Unary;

#define WF2_UNARY(FN, EXPR) \
  template <> struct Unary<FN> { __device__ __forceinline__ static float apply(float x) { return (EXPR); } };

WF2_UNARY(WF2CU_SIN, sinf(x))
WF2_UNARY(WF2CU_COS, cosf(x))
WF2_UNARY(WF2CU_TAN, (float)tan((double)x))
WF2_UNARY(WF2CU_ASIN, (float)asin((double)x))
WF2_UNARY(WF2CU_ACOS, (float)acos((double)x))
WF2_UNARY(WF2CU_ATAN, atanf(x))
WF2_UNARY(WF2CU_SINH, (float)sinh((double)x))
WF2_UNARY(WF2CU_COSH, (float)cosh((double)x))
WF2_UNARY(WF2CU_TANH, tanhf(x))
WF2_UNARY(WF2CU_ABS, fabsf(x))
WF2_UNARY(WF2CU_EXP, expf(x))
WF2_UNARY(WF2CU_LOGE, logf(x))
WF2_UNARY(WF2CU_LOG2, log2f(x))
WF2_UNARY(WF2CU_LOG10, (float)log10((double)x))
WF2_UNARY(WF2CU_LOG1P, log1pf(x))
WF2_UNARY(WF2CU_SQRT, __fsqrt_rn(x))
WF2_UNARY(WF2CU_CBRT, cbrtf(x))
// Sign: LispTensor `(signum x)` (`backends/lisp/mathematics.lisp`): -1, 0, +1 (0 stays 0)
WF2_UNARY(WF2CU_SIGN, (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : x))
WF2_UNARY(WF2CU_SQUARE, __fmul_rn(x, x))
WF2_UNARY(WF2CU_NEG, -x)
WF2_UNARY(WF2CU_RECIP, __fdiv_rn(1.0f, x))
#undef WF2_UNARY

// runtime-dispatched version for the fused-program interpreter
__device__ __forceinline__ float unary_dispatch(int fn, float x) {
  switch (fn) {
    case WF2CU_SIN: return Unary<WF2CU_SIN>::apply(x);
    case WF2CU_COS: return Unary<WF2CU_COS>::apply(x);
    case WF2CU_TAN: return Unary<WF2CU_TAN>::apply(x);
    case WF2CU_ASIN: return Unary<WF2CU_ASIN>::apply(x);
    case WF2CU_ACOS: return Unary<WF2CU_ACOS>::apply(x);
    case WF2CU_ATAN: return Unary<WF2CU_ATAN>::apply(x);
    case WF2CU_SINH: return Unary<WF2CU_SINH>::apply(x);
    case WF2CU_COSH: return Unary<WF2CU_COSH>::apply(x);
    case WF2CU_TANH: return Unary<WF2CU_TANH>::apply(x);
    case WF2CU_ABS: return Unary<WF2CU_ABS>::apply(x);
    case WF2CU_EXP: return Unary<WF2CU_EXP>::apply(x);
    case WF2CU_LOGE: return Unary<WF2CU_LOGE>::apply(x);
    case WF2CU_LOG2: return Unary<WF2CU_LOG2>::apply(x);
    case WF2CU_LOG10: return Unary<WF2CU_LOG10>::apply(x);
    case WF2CU_LOG1P: return Unary<WF2CU_LOG1P>::apply(x);
    case WF2CU_SQRT: return Unary<WF2CU_SQRT>::apply(x);
    case WF2CU_CBRT: return Unary<WF2CU_CBRT>::apply(x);
    case WF2CU_SIGN: return Unary<WF2CU_SIGN>::apply(x);
    case WF2CU_SQUARE: return Unary<WF2CU_SQUARE>::apply(x);
    case WF2CU_NEG: return Unary<WF2CU_NEG>::apply(x);
    case WF2CU_RECIP: return Unary<WF2CU_RECIP>::apply(x);
    default: return x;
  }
}

// powf: CUDA documents powf above 1 ulp; the reference is SLEEF powf u10 (body) / libm pow in
// double (tail).  Use the double route.
// `waffe2_spow` is SLEEF powf (u10).  The general case goes through double-precision pow (correctly rounded to
// fp32 in practice).  The exponents the models actually use — 2 in LayerNorm / MSE, 3 in GELU
// (examples/gpt-2/utils.lisp:114-137) — are exact products in double (x^2: 48 bits, x^3: one double rounding), so
// they give the same fp32 result as the pow route at a fraction of its cost (58 us -> HBM speed on (8 1024 768)).
__device__ __forceinline__ float pow_ref(float x, float n) {
  if (n == 2.0f) return __fmul_rn(x, x);
  if (n == 3.0f) { const double d = (double)x; return (float)(d * d * d); }
  return (float)pow((double)x, (double)n);
}

// `!gelu` as cl-waffe2 writes it (nn/activation.lisp:61-66; examples/gpt-2/utils.lisp:129-137): 0.5 x (1 + tanh(c2 (x + c1 x^3))),
// with exactly the operations and order of the fused program PGelu (fused.cu) / ops.gelu, so a GEMM epilogue that applies it
// gives the same bits as the separate node.  c1 = 0.044715, c2 = (coerce (sqrt (/ 2.0 pi)) 'single-float).
__device__ __forceinline__ float gelu_tanh_ref(float x) {
  float t = __fmul_rn(__fmul_rn(x, x), x);
  t = __fadd_rn(x, __fmul_rn(t, 0.044715f));
  t = Unary<WF2CU_TANH>::apply(__fmul_rn(0.7978845608028654f, t));
  t = __fadd_rn(1.0f, t);
  return __fmul_rn(__fmul_rn(0.5f, x), t);
}

template <int OP> struct Binary;
template <> struct Binary<WF2CU_ADD> { __device__ __forceinline__ static float apply(float a, float b) { return __fadd_rn(a, b); } };
template <> struct Binary<WF2CU_SUB> { __device__ __forceinline__ static float apply(float a, float b) { return __fsub_rn(a, b); } };
template <> struct Binary<WF2CU_MUL> { __device__ __forceinline__ static float apply(float a, float b) { return __fmul_rn(a, b); } };
template <> struct Binary<WF2CU_DIV> { __device__ __forceinline__ static float apply(float a, float b) { return __fdiv_rn(a, b); } };
template <> struct Binary<WF2CU_COPY> { __device__ __forceinline__ static float apply(float a, float b) { return b; } };

__device__ __forceinline__ bool cmp_dispatch(int op, float a, float b) {
  // ordered, quiet compares like the reference's _CMP_*_OQ (`avx2.h:178-244`): NaN -> false
  switch (op) {
    case WF2CU_EQ: return a == b;
    case WF2CU_LT: return a < b;
    case WF2CU_LE: return a <= b;
    case WF2CU_GT: return a > b;
    default: return a >= b;
  }
}

}  // namespace wf2
