// misc.cu — max/min/argmax rows, fused optimizer steps and the GPT-2 embedding gather
// (the fused element-wise chains live in fused.cu).
#include "common.cuh"
#include "math.cuh"

namespace wf2 {

// ------------------------------------------------------------------ max / min / argmax / argmin
// MaxValue/MinValue: waffe2_s{max,min} per row (`backends/cpu/matrix-ops.lisp:332-369`,
// `cl-waffe2-simd.c:316-365`).  ArgMax/ArgMin: isamax_/isamin_ - 1
// (`backends/cpu/matrix-ops.lisp:240-290`): BLAS i?amax ranks |x| and returns the FIRST hit.
struct RowRedArgs {
  const float* x;
  void* out;
  int nouter;
  uint32_t oshape[6];
  int64_t ostride_x[6], ostride_o[6];
  int64_t cols, sx;
  uint32_t rows;
};

template <bool IS_MIN, bool ARG>
__global__ void __launch_bounds__(256) rowmaxmin_kernel(const __grid_constant__ RowRedArgs a) {
  const uint32_t row_id = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row_id >= a.rows) return;
  const int lane = threadIdx.x & 31;
  uint32_t row = row_id;
  int64_t xo = 0, oo = 0;
  for (int d = a.nouter - 1; d >= 0; --d) {
    const uint32_t q = row / a.oshape[d], r = row - q * a.oshape[d];
    row = q; xo += (int64_t)r * a.ostride_x[d]; oo += (int64_t)r * a.ostride_o[d];
  }
  const float* p = a.x + xo;
  // key = value (max/min) or |value| (arg variants)
  float best = 0.f; int64_t bi = -1;
  for (int64_t c = lane; c < a.cols; c += 32) {
    float v = p[c * a.sx];
    if (ARG) v = fabsf(v);
    const bool better = (bi < 0) || (IS_MIN ? (v < best) : (v > best));
    if (better) { best = v; bi = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const long long oi = __shfl_xor_sync(0xffffffffu, (long long)bi, o);
    bool take = false;
    if (oi >= 0) {
      if (bi < 0) take = true;
      else if (IS_MIN ? (ov < best) : (ov > best)) take = true;
      else if (ov == best && oi < bi) take = true;   // first occurrence wins
    }
    if (take) { best = ov; bi = oi; }
  }
  if (lane == 0) {
    if (ARG) reinterpret_cast<uint32_t*>(a.out)[oo] = (uint32_t)bi;
    else reinterpret_cast<float*>(a.out)[oo] = best;
  }
}

static int launch_rowmaxmin(bool is_min, bool arg, const wf2cu_view* x, const wf2cu_view* out, const char* who) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(x && out && x->base && out->base, "%s: null view", who);
  WF2_CHECK(x->rank == out->rank && x->rank >= 1 && x->rank <= 7, "%s: bad rank", who);
  const int last = x->rank - 1;
  RowRedArgs a;
  memset(&a, 0, sizeof(a));
  a.x = fptr(x);
  a.out = arg ? (void*)(reinterpret_cast<uint32_t*>(out->base) + out->offset) : (void*)fptr(out);
  int64_t rows = 1;
  for (int i = 0; i < last; ++i) {
    WF2_CHECK(x->shape[i] == out->shape[i], "%s: shape mismatch on axis %d", who, i);
    a.oshape[i] = (uint32_t)x->shape[i]; a.ostride_x[i] = x->stride[i]; a.ostride_o[i] = out->stride[i];
    rows *= x->shape[i];
  }
  a.nouter = last;
  a.cols = x->shape[last]; a.sx = x->stride[last];
  WF2_CHECK(a.cols >= 1, "%s: empty reduction axis", who);
  if (rows == 0) return 0;
  a.rows = (uint32_t)rows;
  const unsigned grid = (unsigned)((rows + 7) / 8);
  cudaStream_t st = S().stream;
  if (arg) {
    if (is_min) rowmaxmin_kernel<true, true><<<grid, 256, 0, st>>>(a);
    else rowmaxmin_kernel<false, true><<<grid, 256, 0, st>>>(a);
  } else {
    if (is_min) rowmaxmin_kernel<true, false><<<grid, 256, 0, st>>>(a);
    else rowmaxmin_kernel<false, false><<<grid, 256, 0, st>>>(a);
  }
  return post_launch(who);
}

// ------------------------------------------------------------------ optimizers
// SGD: `(A-=B param (A*=scal grad lr))` (`optimizers/impls/sgd.lisp:20-26`): p <- p - (g * lr)
__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ p, const float* __restrict__ g, int64_t n, float lr) {
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride)
    p[i] = __fsub_rn(p[i], __fmul_rn(g[i], lr));
}

// Adam exactly as the reference composes it (`optimizers/impls/adam.lisp:45-67,96-116`):
//   Adam-Step-M     : m <- m + (g - m) * (1 - b1)
//   Adam-Step-V     : v <- v + (1 - b2) * (g*g - v)
//   Adam-Step-Param : p <- p - (m * lr_t) / (eps + sqrt(v))
// lr_t = lr * sqrt(1 - b2^n) / (1 - b1^n) is computed on the host (`adam-step-lr`, :81-94); it is
// read from device memory when lr_dev != NULL so a captured CUDA graph can be replayed with a new
// value every step.  Three `defmodel-as` composites per parameter (9 memory passes) become one
// kernel over the whole flat parameter bucket.
__device__ __forceinline__ float adam_one(float gi, float gscale, float& m, float& v, float p, float lr_t, float omb1, float omb2, float eps) {
  if (gscale != 1.0f) gi = __fmul_rn(gi, gscale);
  const float mi = __fadd_rn(m, __fmul_rn(__fsub_rn(gi, m), omb1));
  const float vi = __fadd_rn(v, __fmul_rn(omb2, __fsub_rn(__fmul_rn(gi, gi), v)));
  m = mi; v = vi;
  return __fsub_rn(p, __fdiv_rn(__fmul_rn(mi, lr_t), __fadd_rn(eps, __fsqrt_rn(vi))));
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                   float lr_t, const float* __restrict__ lr_dev, float b1, float b2,
                                                   float eps, float gscale) {
  if (lr_dev) lr_t = *lr_dev;
  const float omb1 = __fsub_rn(1.0f, b1), omb2 = __fsub_rn(1.0f, b2);
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    float mi = m[i], vi = v[i];
    p[i] = adam_one(g[i], gscale, mi, vi, p[i], lr_t, omb1, omb2, eps);
    m[i] = mi; v[i] = vi;
  }
}

// The same update with 128-bit accesses: 7 streams (4 read, 3 written), two vectors per stream in flight per thread.
// Element-wise, so the result is bit-identical to adam_kernel.
__global__ void __launch_bounds__(256) adam_vec_kernel(float4* __restrict__ p, const float4* __restrict__ g,
                                                       float4* __restrict__ m, float4* __restrict__ v, uint32_t n4,
                                                       float lr_t, const float* __restrict__ lr_dev, float b1, float b2,
                                                       float eps, float gscale) {
  pdl_prologue();
  if (lr_dev) lr_t = *lr_dev;
  const float omb1 = __fsub_rn(1.0f, b1), omb2 = __fsub_rn(1.0f, b2);
  const uint32_t i0 = blockIdx.x * 512u + threadIdx.x, i1 = i0 + 256u;
  const bool ok0 = i0 < n4, ok1 = i1 < n4;
  float4 P0, G0, M0, V0, P1, G1, M1, V1;
  if (ok0) { G0 = g[i0]; M0 = m[i0]; V0 = v[i0]; P0 = p[i0]; }
  if (ok1) { G1 = g[i1]; M1 = m[i1]; V1 = v[i1]; P1 = p[i1]; }
  if (ok0) {
    P0.x = adam_one(G0.x, gscale, M0.x, V0.x, P0.x, lr_t, omb1, omb2, eps);
    P0.y = adam_one(G0.y, gscale, M0.y, V0.y, P0.y, lr_t, omb1, omb2, eps);
    P0.z = adam_one(G0.z, gscale, M0.z, V0.z, P0.z, lr_t, omb1, omb2, eps);
    P0.w = adam_one(G0.w, gscale, M0.w, V0.w, P0.w, lr_t, omb1, omb2, eps);
    m[i0] = M0; v[i0] = V0; p[i0] = P0;
  }
  if (ok1) {
    P1.x = adam_one(G1.x, gscale, M1.x, V1.x, P1.x, lr_t, omb1, omb2, eps);
    P1.y = adam_one(G1.y, gscale, M1.y, V1.y, P1.y, lr_t, omb1, omb2, eps);
    P1.z = adam_one(G1.z, gscale, M1.z, V1.z, P1.z, lr_t, omb1, omb2, eps);
    P1.w = adam_one(G1.w, gscale, M1.w, V1.w, P1.w, lr_t, omb1, omb2, eps);
    m[i1] = M1; v[i1] = V1; p[i1] = P1;
  }
}

// ------------------------------------------------------------------ GPT-2 embedding
// out[b,p,:] = wte[round(ctx[b,p]),:] + wpe[p,:]   (`examples/gpt-2/utils.lisp:19-112`)
__global__ void __launch_bounds__(256) embedding_kernel(const float* __restrict__ ctx, int64_t seq,
                                                        const float* __restrict__ wte, const float* __restrict__ wpe,
                                                        int64_t dim, float* __restrict__ out) {
  const int64_t bp = blockIdx.x;
  const int64_t pos = bp % seq;
  const int64_t tok = (int64_t)rintf(ctx[bp]);
  const float* e = wte + tok * dim;
  const float* w = wpe + pos * dim;
  float* o = out + bp * dim;
  for (int64_t d = threadIdx.x; d < dim; d += 256) o[d] = __fadd_rn(e[d], w[d]);
}

}  // namespace wf2

using namespace wf2;

extern "C" {

int wf2cu_reduce_maxmin(int is_min, const wf2cu_view* x, const wf2cu_view* out) {
  return launch_rowmaxmin(is_min != 0, false, x, out, "wf2cu_reduce_maxmin");
}

int wf2cu_argmaxmin(int is_min, const wf2cu_view* x, const wf2cu_view* out_u32) {
  return launch_rowmaxmin(is_min != 0, true, x, out_u32, "wf2cu_argmaxmin");
}

int wf2cu_sgd_step(void* param, int64_t p_off, const void* grad, int64_t g_off, int64_t n, float lr) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(param && grad && n >= 0, "wf2cu_sgd_step: bad arguments");
  if (n == 0) return 0;
  int64_t grid = (n + 255) / 256;
  const int64_t cap = (int64_t)S().sm_count * 8;
  if (grid > cap) grid = cap;
  sgd_kernel<<<(unsigned)grid, 256, 0, S().stream>>>((float*)param + p_off, (const float*)grad + g_off, n, lr);
  return post_launch("wf2cu_sgd_step");
}

int wf2cu_adam_step(void* param, int64_t p_off, const void* grad, int64_t g_off, void* m, int64_t m_off,
                    void* v, int64_t v_off, int64_t n, float lr_t, const void* lr_t_dev, float beta1, float beta2,
                    float eps, float grad_scale) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(param && grad && m && v && n >= 0, "wf2cu_adam_step: bad arguments");
  if (n == 0) return 0;
  float* pp = (float*)param + p_off; const float* gp = (const float*)grad + g_off;
  float* mp = (float*)m + m_off; float* vp = (float*)v + v_off;
  const auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if ((n & 3) == 0 && n / 4 < (int64_t(1) << 31) && al16(pp) && al16(gp) && al16(mp) && al16(vp)) {
    const uint32_t n4 = (uint32_t)(n / 4);
    WF2_CUDA(launch_pdl(adam_vec_kernel, dim3((n4 + 511u) / 512u), dim3(256), 0, S().stream, (float4*)pp, (const float4*)gp, (float4*)mp,
                        (float4*)vp, n4, lr_t, (const float*)lr_t_dev, beta1, beta2, eps, grad_scale));
    return post_launch("wf2cu_adam_step(vec)");
  }
  int64_t grid = (n + 255) / 256;
  const int64_t cap = (int64_t)S().sm_count * 8;
  if (grid > cap) grid = cap;
  adam_kernel<<<(unsigned)grid, 256, 0, S().stream>>>((float*)param + p_off, (const float*)grad + g_off,
                                                     (float*)m + m_off, (float*)v + v_off, n, lr_t,
                                                     (const float*)lr_t_dev, beta1, beta2, eps, grad_scale);
  return post_launch("wf2cu_adam_step");
}

int wf2cu_embedding_gather(const void* ctx, int64_t batch, int64_t seq, const void* wte, const void* wpe,
                           int64_t dim, void* out) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(ctx && wte && wpe && out, "wf2cu_embedding_gather: null pointer");
  if (batch * seq == 0) return 0;
  embedding_kernel<<<(unsigned)(batch * seq), 256, 0, S().stream>>>((const float*)ctx, seq, (const float*)wte,
                                                                   (const float*)wpe, dim, (float*)out);
  return post_launch("wf2cu_embedding_gather");
}

}  // extern "C"
