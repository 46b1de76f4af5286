// misc.cu — max/min/argmax rows, fused optimizer steps, the GPT-2 embedding gather and the
// fused element-wise program interpreter.
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
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                   float lr_t, const float* __restrict__ lr_dev, float b1, float b2,
                                                   float eps, float gscale) {
  if (lr_dev) lr_t = *lr_dev;
  const float omb1 = __fsub_rn(1.0f, b1), omb2 = __fsub_rn(1.0f, b2);
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    float gi = g[i];
    if (gscale != 1.0f) gi = __fmul_rn(gi, gscale);
    const float m0 = m[i], v0 = v[i];
    const float mi = __fadd_rn(m0, __fmul_rn(__fsub_rn(gi, m0), omb1));
    const float vi = __fadd_rn(v0, __fmul_rn(omb2, __fsub_rn(__fmul_rn(gi, gi), v0)));
    m[i] = mi; v[i] = vi;
    const float denom = __fadd_rn(eps, __fsqrt_rn(vi));
    p[i] = __fsub_rn(p[i], __fdiv_rn(__fmul_rn(mi, lr_t), denom));
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

// ------------------------------------------------------------------ fused program interpreter
constexpr int kMaxInstr = 64;
constexpr int kMaxRegs = 16;
struct FusedArgs {
  int n_instr;
  int nviews;
  int4 prog[kMaxInstr];
  float* base[kMaxOperands];
  int rank;
  int64_t shape[kMaxRank];
  int64_t stride[kMaxOperands][kMaxRank];
  int64_t numel;
  int vec;  // all views contiguous (stride 1 or 0 over a rank-1 plan) and 16-byte aligned
};

__device__ __forceinline__ float fused_eval(const int4 ins, const float* r) {
  const int opc = ins.x & 0xff;
  switch (opc) {
    case WF2CU_P_IMM: return __int_as_float(ins.w);
    case WF2CU_P_ADD: return __fadd_rn(r[ins.z], r[ins.w]);
    case WF2CU_P_SUB: return __fsub_rn(r[ins.z], r[ins.w]);
    case WF2CU_P_MUL: return __fmul_rn(r[ins.z], r[ins.w]);
    case WF2CU_P_DIV: return __fdiv_rn(r[ins.z], r[ins.w]);
    case WF2CU_P_UNARY: return unary_dispatch(ins.w, r[ins.z]);
    case WF2CU_P_POW: return pow_ref(r[ins.z], r[ins.w]);
    case WF2CU_P_CMPSEL: return cmp_dispatch(ins.x >> 8, r[ins.z], r[ins.w]) ? 1.0f : 0.0f;
    case WF2CU_P_MOV: return r[ins.z];
    default: return 0.f;
  }
}

__global__ void __launch_bounds__(256) fused_kernel(const __grid_constant__ FusedArgs a) {
  const int64_t gstride = (int64_t)gridDim.x * 256;
  if (a.vec) {
    const int64_t n4 = a.numel >> 2;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += gstride) {
      float r[4][kMaxRegs];
      for (int k = 0; k < a.n_instr; ++k) {
        const int4 ins = a.prog[k];
        const int opc = ins.x & 0xff;
        if (opc == WF2CU_P_LOAD) {
          if (a.stride[ins.z][0] == 0) {
            const float s = a.base[ins.z][0];
            r[0][ins.y] = s; r[1][ins.y] = s; r[2][ins.y] = s; r[3][ins.y] = s;
          } else {
            const float4 v = reinterpret_cast<const float4*>(a.base[ins.z])[i];
            r[0][ins.y] = v.x; r[1][ins.y] = v.y; r[2][ins.y] = v.z; r[3][ins.y] = v.w;
          }
        } else if (opc == WF2CU_P_STORE) {
          reinterpret_cast<float4*>(a.base[ins.z])[i] = make_float4(r[0][ins.y], r[1][ins.y], r[2][ins.y], r[3][ins.y]);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) r[e][ins.y] = fused_eval(ins, r[e]);
        }
      }
    }
    return;
  }
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < a.numel; idx += gstride) {
    int64_t off[kMaxOperands];
    for (int j = 0; j < a.nviews; ++j) off[j] = 0;
    int64_t rem = idx;
    for (int d = a.rank - 1; d >= 0; --d) {
      const int64_t q = rem / a.shape[d], rr = rem - q * a.shape[d];
      rem = q;
      for (int j = 0; j < a.nviews; ++j) off[j] += rr * a.stride[j][d];
    }
    float r[kMaxRegs];
    for (int k = 0; k < a.n_instr; ++k) {
      const int4 ins = a.prog[k];
      const int opc = ins.x & 0xff;
      if (opc == WF2CU_P_LOAD) r[ins.y] = a.base[ins.z][off[ins.z]];
      else if (opc == WF2CU_P_STORE) a.base[ins.z][off[ins.z]] = r[ins.y];
      else r[ins.y] = fused_eval(ins, r);
    }
  }
}

// ---- compiled form of one well-known program: `!gelu` as cl-waffe2 writes it (nn/activation.lisp:61-66;
// examples/gpt-2/utils.lisp:129-137): 0.5 x (1 + tanh(c2 (x + c1 x^3))), same operation order and roundings as
// the interpreted program.  The interpreter keeps its register file in local memory (runtime-indexed) and
// reached 0.94 TB/s on GPT-2's (8 1024 3072) activation; this kernel is HBM-bound.
static const int kGeluOps[15][4] = {
    {WF2CU_P_LOAD, 0, 0, 0},  {WF2CU_P_MUL, 1, 0, 0},  {WF2CU_P_MUL, 1, 1, 0},  {WF2CU_P_IMM, 2, 0, -1}, {WF2CU_P_MUL, 1, 1, 2},
    {WF2CU_P_ADD, 1, 0, 1},   {WF2CU_P_IMM, 2, 0, -1}, {WF2CU_P_MUL, 1, 2, 1},  {WF2CU_P_UNARY, 1, 1, WF2CU_TANH}, {WF2CU_P_IMM, 2, 0, -1},
    {WF2CU_P_ADD, 1, 2, 1},   {WF2CU_P_IMM, 2, 0, -1}, {WF2CU_P_MUL, 2, 2, 0},  {WF2CU_P_MUL, 1, 2, 1},  {WF2CU_P_STORE, 1, 1, 0}};

static bool match_gelu(const int32_t* program, int n_instr, float imm[4]) {
  if (n_instr != 15) return false;
  int ni = 0;
  for (int k = 0; k < 15; ++k) {
    const int32_t* q = program + 4 * k;
    if (q[0] != kGeluOps[k][0] || q[1] != kGeluOps[k][1] || q[2] != kGeluOps[k][2]) return false;
    if (kGeluOps[k][0] == WF2CU_P_IMM) { memcpy(&imm[ni++], &q[3], 4); continue; }
    if (kGeluOps[k][0] == WF2CU_P_LOAD || kGeluOps[k][0] == WF2CU_P_STORE) continue;
    if (q[3] != kGeluOps[k][3]) return false;
  }
  return ni == 4;
}

__device__ __forceinline__ float gelu_one(float x, float c1, float c2, float one, float half) {
  float t = __fmul_rn(__fmul_rn(x, x), x);
  t = __fadd_rn(x, __fmul_rn(t, c1));
  t = Unary<WF2CU_TANH>::apply(__fmul_rn(c2, t));
  t = __fadd_rn(one, t);
  return __fmul_rn(__fmul_rn(half, x), t);
}

__global__ void __launch_bounds__(256) gelu_vec_kernel(const float4* x, float4* out, int64_t n4,   // may alias (in-place gelu)
                                                       float c1, float c2, float one, float half) {
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += stride) {
    const float4 v = x[i];
    out[i] = make_float4(gelu_one(v.x, c1, c2, one, half), gelu_one(v.y, c1, c2, one, half),
                         gelu_one(v.z, c1, c2, one, half), gelu_one(v.w, c1, c2, one, half));
  }
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

int wf2cu_fused(const int32_t* program, int n_instr, const wf2cu_view* views, int nviews) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(program && views, "wf2cu_fused: null argument");
  WF2_CHECK(n_instr >= 1 && n_instr <= kMaxInstr, "wf2cu_fused: %d instructions (max %d)", n_instr, kMaxInstr);
  WF2_CHECK(nviews >= 1 && nviews <= kMaxOperands, "wf2cu_fused: %d views (max %d)", nviews, kMaxOperands);
  const wf2cu_view* vp[kMaxOperands];
  for (int j = 0; j < nviews; ++j) vp[j] = &views[j];
  Plan p;
  int rc = build_plan(vp, nviews, &p, "wf2cu_fused");
  if (rc) return rc;
  if (p.numel == 0) return 0;
  FusedArgs a;
  memset(&a, 0, sizeof(a));
  a.n_instr = n_instr; a.nviews = nviews;
  bool stored[kMaxOperands] = {false};
  for (int k = 0; k < n_instr; ++k) {
    const int32_t* q = program + 4 * k;
    const int opc = q[0] & 0xff;
    WF2_CHECK(opc >= WF2CU_P_LOAD && opc <= WF2CU_P_MOV, "wf2cu_fused: bad opcode %d at %d", opc, k);
    WF2_CHECK(q[1] >= 0 && q[1] < kMaxRegs, "wf2cu_fused: bad dst register at %d", k);
    if (opc == WF2CU_P_LOAD || opc == WF2CU_P_STORE) {
      WF2_CHECK(q[2] >= 0 && q[2] < nviews, "wf2cu_fused: bad view index at %d", k);
      if (opc == WF2CU_P_STORE) stored[q[2]] = true;
    } else if (opc != WF2CU_P_IMM) {
      WF2_CHECK(q[2] >= 0 && q[2] < kMaxRegs, "wf2cu_fused: bad source register at %d", k);
      if (opc != WF2CU_P_UNARY && opc != WF2CU_P_MOV)
        WF2_CHECK(q[3] >= 0 && q[3] < kMaxRegs, "wf2cu_fused: bad source register at %d", k);
    }
    a.prog[k] = make_int4(q[0], q[1], q[2], q[3]);
  }
  bool vec = (p.rank == 1) && (p.numel % 4 == 0);
  for (int j = 0; j < nviews; ++j) {
    a.base[j] = fptr(&views[j]);
    for (int d = 0; d < p.rank; ++d) {
      a.stride[j][d] = p.stride[j][d];
      if (stored[j] && p.stride[j][d] == 0 && p.shape[d] > 1)
        return set_error(WF2CU_ERR_INVALID, "wf2cu_fused: stored view %d has a broadcast axis", j);
    }
    const int64_t s0 = p.stride[j][0];
    if (!(s0 == 1 || s0 == 0)) vec = false;
    if (s0 == 1 && (reinterpret_cast<uintptr_t>(a.base[j]) & 15)) vec = false;
  }
  a.rank = p.rank;
  for (int d = 0; d < p.rank; ++d) a.shape[d] = p.shape[d];
  a.numel = p.numel;
  a.vec = vec ? 1 : 0;
  int64_t work = vec ? p.numel / 4 : p.numel;
  int64_t grid = (work + 255) / 256;
  const int64_t cap = (int64_t)S().sm_count * 8;
  if (grid > cap) grid = cap;
  float imm[4];
  if (vec && nviews == 2 && p.stride[0][0] == 1 && p.stride[1][0] == 1 && match_gelu(program, n_instr, imm)) {
    int64_t g = (work + 255) / 256;
    const int64_t gcap = (int64_t)S().sm_count * 32;
    if (g > gcap) g = gcap;
    gelu_vec_kernel<<<(unsigned)g, 256, 0, S().stream>>>(reinterpret_cast<const float4*>(a.base[0]), reinterpret_cast<float4*>(a.base[1]),
                                                         work, imm[0], imm[1], imm[2], imm[3]);
    return post_launch("wf2cu_fused(gelu)");
  }
  fused_kernel<<<(unsigned)grid, 256, 0, S().stream>>>(a);
  return post_launch("wf2cu_fused");
}

}  // extern "C"
