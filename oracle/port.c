/*
 * oracle/port.c — TEST INFRASTRUCTURE, NOT PRODUCT.  Never linked into libwaffe2cuda.so.
 *
 * A scalar C restatement of the leaf kernels of the reference's cl-waffe2-simd extension, with
 * the same C ABI, so the loop-nest driver (oracle/loopnest.py) can run with or without the real
 * reference build (oracle/_ref/libwaffe2simd_ref.so).  Follows
 *   /root/reference/cl-waffe2-simd/kernels/cl-waffe2-simd.c
 *     :153-213  define_arithmetic_func   x[i*incx] <- x[i*incx] OP y[i*incy]
 *     :316-347  define_maxmin            y[0] <- max/min_i x[i*incx]
 *     :385-424  define_cmp               out <- (x CMP y) ? then : else
 *     :484-522  define_scmp              out <- (x CMP scal) ? then : else
 *     :586-642  define_math_func         y[i*incy] <- f(x[i*incx])
 *     :708-765  define_math_func_pow     y <- x ^ n
 *
 * Differences from the reference build, all rounding-neutral or documented:
 *  - add/sub/mul/div/compare/max/min are exactly rounded IEEE operations, so scalar and AVX2
 *    evaluation give identical bits (compiled with -ffp-contract=off).  Sequential order is kept,
 *    which matters when incx == 0 (the reference skips its SIMD body for incx < 1, :159, so a
 *    stride-0 destination accumulates strictly left to right — this is how `!sum` runs).
 *  - unary math uses the reference's TAIL arithmetic for every element: double-precision libm
 *    narrowed to float (:636-641 with e.g. define_math_func(ssin, ..., sin) :645).  The reference's
 *    SIMD body uses SLEEF u10 instead; both are within 1 ulp of the true value, so they may differ
 *    from each other by an ulp.  Use _ref for the body arithmetic.
 *  - asinh/acosh/atanh are not provided: the reference's SIMD body for them is wrong (:672-679
 *    call sinh/cosh/tanh) and no CPUTensor node reaches them.
 */
#include <math.h>
#include <stdint.h>

#define ARITH(NAME, EXPR)                                                              \
  void waffe2_s##NAME(const long n, float* x, const long incx, float* y, const long incy) { \
    for (long i = 0; i < n; ++i) {                                                     \
      float a = x[0], b = y[0];                                                        \
      (void)a; (void)b;                                                                \
      x[0] = (EXPR);                                                                   \
      x += incx;                                                                       \
      y += incy;                                                                       \
    }                                                                                  \
  }
ARITH(add, a + b)
ARITH(sub, a - b)
ARITH(mul, a * b)
ARITH(div, a / b)
/* MoveTensorNode goes through BLAS scopy_ for floats (backends/cpu/arithmetic.lisp:39-61);
 * same contract, provided here under the waffe2 naming for the driver. */
ARITH(copy, b)

#define MATH(NAME, FN)                                                                 \
  void waffe2_s##NAME(const long n, float* x, const long incx, float* y, const long incy) { \
    for (long i = 0; i < n; ++i) {                                                     \
      y[0] = (float)FN((double)x[0]);                                                  \
      x += incx;                                                                       \
      y += incy;                                                                       \
    }                                                                                  \
  }
MATH(sin, sin)
MATH(cos, cos)
MATH(tan, tan)
MATH(asin, asin)
MATH(acos, acos)
MATH(atan, atan)
MATH(sinh, sinh)
MATH(cosh, cosh)
MATH(tanh, tanh)
MATH(log, log)
MATH(log1p, log1p)
MATH(log2, log2)
MATH(log10, log10)
MATH(exp, exp)
MATH(abs, fabs)
MATH(sqrt, sqrt)
MATH(cbrt, cbrt)

void waffe2_spow(const long n, float* x, const long incx, float pow_n, float* y, const long incy) {
  for (long i = 0; i < n; ++i) {
    y[0] = (float)pow((double)x[0], (double)pow_n);
    x += incx;
    y += incy;
  }
}

void waffe2_smax(const long n, float* x, const long incx, float* y) {
  float m = x[0];
  for (long i = 0; i < n; ++i) { m = (m > x[0]) ? m : x[0]; x += incx; }
  y[0] = m;
}
void waffe2_smin(const long n, float* x, const long incx, float* y) {
  float m = x[0];
  for (long i = 0; i < n; ++i) { m = (m < x[0]) ? m : x[0]; x += incx; }
  y[0] = m;
}

#define CMP(NAME, OP)                                                                   \
  void waffe2_s##NAME(const long n, float* x, const long incx, float* y, const long incy, \
                      float* out, const long inco, float then_value, float else_value) { \
    for (long i = 0; i < n; ++i) {                                                      \
      out[0] = (x[0] OP y[0]) ? then_value : else_value;                                \
      x += incx; y += incy; out += inco;                                                \
    }                                                                                   \
  }                                                                                     \
  void waffe2_s##NAME##_scal(const long n, float* x, const long incx, float* out,       \
                             const long inco, float y, float then_value, float else_value) { \
    for (long i = 0; i < n; ++i) {                                                      \
      out[0] = (x[0] OP y) ? then_value : else_value;                                   \
      x += incx; out += inco;                                                           \
    }                                                                                   \
  }
CMP(eq, ==)
CMP(lt, <)
CMP(le, <=)
CMP(gt, >)
CMP(ge, >=)

/* OpenBLAS stand-ins used only when scipy's BLAS is not importable: isamax_/isamin_ semantics
 * (`backends/cpu/matrix-ops.lisp:240-290`: 1-based index of max |x|, first occurrence). */
long wf2port_isamax(const long n, const float* x, const long incx) {
  long best = 0; float bv = fabsf(x[0]);
  for (long i = 1; i < n; ++i) { float v = fabsf(x[i * incx]); if (v > bv) { bv = v; best = i; } }
  return best + 1;
}
long wf2port_isamin(const long n, const float* x, const long incx) {
  long best = 0; float bv = fabsf(x[0]);
  for (long i = 1; i < n; ++i) { float v = fabsf(x[i * incx]); if (v < bv) { bv = v; best = i; } }
  return best + 1;
}

/* Plain triple-loop fp32 GEMM, row-major, used only as a fall-back checker when no BLAS is
 * importable.  C[m,n] = sum_k opA[m,k] * opB[k,n], accumulating in float in k order. */
void wf2port_sgemm(int ta, int tb, long m, long n, long k, const float* A, long lda,
                   const float* B, long ldb, float* C, long ldc) {
  for (long i = 0; i < m; ++i)
    for (long j = 0; j < n; ++j) {
      float acc = 0.f;
      for (long p = 0; p < k; ++p) {
        float a = ta ? A[p * lda + i] : A[i * lda + p];
        float b = tb ? B[j * ldb + p] : B[p * ldb + j];
        acc += a * b;
      }
      C[i * ldc + j] = acc;
    }
}

/* Unfold leaves (`cl-waffe2-simd/kernels/im2col.c:63-115,395-448`), restated for `:order :column`
 * (C order, the default).  The image handed in is already padded (`nn/im2col.lisp:88-98`); for kernel
 * offset (kh, kw) the reference visits image rows y = kh, kh+stride_h, ... while
 * y < dilation_h*kh - pad_h + stride_h*H_out (unsigned arithmetic, so "- pad_h" wraps and comes back),
 * and likewise for columns.  Elements of `col` outside that range are left untouched.
 * The `_row` variants (`im2col.c:227-279,562-615`) index `col` with strides that are not a bijection
 * (the W_out index is multiplied by C*K_H*K_W*H_out*W_out) and write out of bounds for W_out > N:
 * they are not restated and the CUDA backend declines `:order :row` for these two nodes. */
#include <stdint.h>
#define WF2PORT_UNFOLD_ARGS                                                                           \
  const uint64_t N, const uint64_t C, const uint64_t H, const uint64_t W, const uint64_t Ho,           \
  const uint64_t Wo, const uint64_t KH, const uint64_t KW, const uint64_t pad_h, const uint64_t pad_w, \
  const uint64_t sh, const uint64_t sw, const uint64_t dh, const uint64_t dw

static void unfold_walk(int scatter, float* col, WF2PORT_UNFOLD_ARGS, float* im) {
  if (scatter) for (uint64_t i = 0; i < N * C * H * W; ++i) im[i] = 0.0f;
  for (uint64_t n = 0; n < N; ++n)
    for (uint64_t kh = 0; kh < KH; ++kh) {
      const uint64_t y_end = dh * kh - pad_h + sh * Ho;
      for (uint64_t kw = 0; kw < KW; ++kw) {
        const uint64_t x_end = dw * kw - pad_w + sw * Wo;
        uint64_t j = 0;
        for (uint64_t y = kh; y < y_end; y += sh, ++j) {
          uint64_t i = 0;
          for (uint64_t x = kw; x < x_end; x += sw, ++i)
            for (uint64_t c = 0; c < C; ++c) {
              float* pc = col + ((((n * C + c) * KH + kh) * KW + kw) * Ho + j) * Wo + i;
              float* pi = im + ((n * C + c) * H + y) * W + x;
              if (scatter) *pi += *pc; else *pc = *pi;
            }
        }
      }
    }
}

void waffe2_im2col_s_column(float* col, WF2PORT_UNFOLD_ARGS, const float* im) {
  unfold_walk(0, col, N, C, H, W, Ho, Wo, KH, KW, pad_h, pad_w, sh, sw, dh, dw, (float*)im);
}
void waffe2_col2im_s_column(const float* col, WF2PORT_UNFOLD_ARGS, float* im) {
  unfold_walk(1, (float*)col, N, C, H, W, Ho, Wo, KH, KW, pad_h, pad_w, sh, sw, dh, dw, im);
}
