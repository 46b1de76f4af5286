// elementwise.cu — the strided / broadcast element-wise loops of cl-waffe2 as whole-view kernels.
//
// The reference lowers an element-wise node to a Lisp loop nest around ONE 1-D foreign call per
// innermost row (`source/vm/generic-tensor/call-with-view.lisp:24-229`,
// `do-compiled-loop.lisp:170-298`) into `waffe2_s{add,sub,mul,div}` / `scopy_` /
// `waffe2_s<math>` (`cl-waffe2-simd/kernels/cl-waffe2-simd.c:153-250,586-768`).  Here the whole
// view (shape, per-operand strides incl. 0 = broadcast and negative = reversed slice, base
// offset) is a kernel argument and one launch covers the tensor.
//
// Kernel variants, picked per call from the merged iteration plan (common.cuh:build_plan):
//   vec   : innermost axis contiguous (stride 1) or broadcast (stride 0) for every operand,
//           16-byte aligned -> 128-bit loads/stores, 4 independent vectors in flight per thread.
//   tr    : one operand is contiguous along a different axis than the output (a `!t` /
//           permuted operand) -> 32x32 tile staged through shared memory, both sides coalesced.
//   gen   : anything else (negative steps, odd alignments, rank <= 8) -> scalar strided.
// All of it is HBM-bound fp32 work: no tensor cores.
#include "common.cuh"
#include "math.cuh"

namespace wf2 {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;

// ---------------------------------------------------------------- functors
template <int OP> struct FBinary {  // NIN = 2
  __device__ __forceinline__ float operator()(const float* in) const { return Binary<OP>::apply(in[0], in[1]); }
};
template <int OP> struct FScalar {  // NIN = 1, dst op scalar
  float s;
  __device__ __forceinline__ float operator()(const float* in) const { return Binary<OP>::apply(in[0], s); }
};
struct FCopy {
  __device__ __forceinline__ float operator()(const float* in) const { return in[0]; }
};
struct FFill {
  float s;
  __device__ __forceinline__ float operator()(const float*) const { return s; }
};
template <int FN> struct FUnary {
  __device__ __forceinline__ float operator()(const float* in) const { return Unary<FN>::apply(in[0]); }
};
struct FPow {
  float n;
  __device__ __forceinline__ float operator()(const float* in) const { return pow_ref(in[0], n); }
};
struct FCmp2 {
  int op; float t, e;
  __device__ __forceinline__ float operator()(const float* in) const { return cmp_dispatch(op, in[0], in[1]) ? t : e; }
};
struct FCmp1 {
  int op; float s, t, e;
  __device__ __forceinline__ float operator()(const float* in) const { return cmp_dispatch(op, in[0], s) ? t : e; }
};

// ---------------------------------------------------------------- kernel arguments
template <int NIN> struct VecArgs {
  float* out;
  const float* in[NIN > 0 ? NIN : 1];
  int outer_rank;                    // 0..4 outer axes (slowest first)
  uint32_t inner4;                   // innermost extent / 4
  uint32_t oshape[4];
  int64_t ostride_out[4];            // element strides of outer axes
  int64_t ostride_in[NIN > 0 ? NIN : 1][4];
  uint32_t inner_bcast;              // bit j: input j has inner stride 0
  int64_t n4;                        // total float4 count
};

template <int NIN> struct GenArgs {
  float* out;
  const float* in[NIN > 0 ? NIN : 1];
  int rank;
  int64_t shape[kMaxRank];
  int64_t stride_out[kMaxRank];
  int64_t stride_in[NIN > 0 ? NIN : 1][kMaxRank];
  int64_t numel;
};

// ---------------------------------------------------------------- vec kernel
template <int NIN, class F>
__global__ void __launch_bounds__(kThreads) ew_vec_kernel(const F f, const __grid_constant__ VecArgs<NIN> a) {
  pdl_prologue();
  // n4 < 2^31 is guaranteed by the dispatcher: all index math below is 32-bit
  const uint32_t base = blockIdx.x * (uint32_t)(kThreads * kUnroll) + threadIdx.x;
  const uint32_t n4 = (uint32_t)a.n4;
  float4 v[NIN > 0 ? NIN : 1][kUnroll];
  int64_t ooff[kUnroll];
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    const uint32_t idx = base + (uint32_t)u * kThreads;
    ooff[u] = -1;
    if (idx < n4) {
      uint32_t c4;
      int64_t o_out = 0;
      int64_t o_in[NIN > 0 ? NIN : 1];
#pragma unroll
      for (int j = 0; j < NIN; ++j) o_in[j] = 0;
      if (a.outer_rank == 0) {
        c4 = idx;
      } else {
        uint32_t row = idx / a.inner4;
        c4 = idx - row * a.inner4;
        for (int d = a.outer_rank - 1; d > 0; --d) {
          const uint32_t q = row / a.oshape[d];
          const uint32_t r = row - q * a.oshape[d];
          row = q;
          o_out += (int64_t)r * a.ostride_out[d];
#pragma unroll
          for (int j = 0; j < NIN; ++j) o_in[j] += (int64_t)r * a.ostride_in[j][d];
        }
        o_out += (int64_t)row * a.ostride_out[0];
#pragma unroll
        for (int j = 0; j < NIN; ++j) o_in[j] += (int64_t)row * a.ostride_in[j][0];
      }
      ooff[u] = o_out + (int64_t)c4 * 4;
#pragma unroll
      for (int j = 0; j < NIN; ++j) {
        if (a.inner_bcast & (1u << j)) {
          const float s = a.in[j][o_in[j]];
          v[j][u] = make_float4(s, s, s, s);
        } else {
          v[j][u] = *reinterpret_cast<const float4*>(a.in[j] + o_in[j] + (int64_t)c4 * 4);
        }
      }
    }
  }
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    if (ooff[u] >= 0) {
      float x0[NIN > 0 ? NIN : 1], x1[NIN > 0 ? NIN : 1], x2[NIN > 0 ? NIN : 1], x3[NIN > 0 ? NIN : 1];
#pragma unroll
      for (int j = 0; j < NIN; ++j) { x0[j] = v[j][u].x; x1[j] = v[j][u].y; x2[j] = v[j][u].z; x3[j] = v[j][u].w; }
      float4 r;
      r.x = f(x0); r.y = f(x1); r.z = f(x2); r.w = f(x3);
      *reinterpret_cast<float4*>(a.out + ooff[u]) = r;
    }
  }
}

// ---------------------------------------------------------------- generic strided kernel
template <int NIN, class F>
__global__ void __launch_bounds__(kThreads) ew_gen_kernel(const F f, const __grid_constant__ GenArgs<NIN> a) {
  pdl_prologue();
  const int64_t gstride = (int64_t)gridDim.x * kThreads;
  for (int64_t idx = (int64_t)blockIdx.x * kThreads + threadIdx.x; idx < a.numel; idx += gstride) {
    int64_t rem = idx, o_out = 0;
    int64_t o_in[NIN > 0 ? NIN : 1];
#pragma unroll
    for (int j = 0; j < NIN; ++j) o_in[j] = 0;
    for (int d = a.rank - 1; d >= 0; --d) {
      const int64_t q = rem / a.shape[d];
      const int64_t r = rem - q * a.shape[d];
      rem = q;
      o_out += r * a.stride_out[d];
#pragma unroll
      for (int j = 0; j < NIN; ++j) o_in[j] += r * a.stride_in[j][d];
    }
    float x[NIN > 0 ? NIN : 1];
#pragma unroll
    for (int j = 0; j < NIN; ++j) x[j] = a.in[j][o_in[j]];
    a.out[o_out] = f(x);
  }
}

// ---------------------------------------------------------------- transposed-operand kernel
// Iteration space: [batch..., J, I] where I is the output's contiguous axis and J is the axis
// along which the "transposed" inputs (mask tr) are contiguous.  Tile = 32(J) x 32(I) x 4 batches
// of rows handled by a 32x8 thread block.
template <int NIN> struct TrArgs {
  float* out;
  const float* in[NIN > 0 ? NIN : 1];
  int nbatch;                       // number of batch axes (0..3)
  uint32_t bshape[3];
  int64_t bstride_out[3];
  int64_t bstride_in[NIN > 0 ? NIN : 1][3];
  int64_t I, J;
  int64_t sI_out, sJ_out;           // sI_out == 1
  int64_t sI_in[NIN > 0 ? NIN : 1], sJ_in[NIN > 0 ? NIN : 1];
  uint32_t tr_mask;                 // bit j: input j is staged through smem (sJ_in == 1)
  uint32_t tilesI, tilesJ;
};

template <int NIN, class F>
__global__ void __launch_bounds__(256) ew_tr_kernel(const F f, const __grid_constant__ TrArgs<NIN> a) {
  __shared__ float tile[NIN > 0 ? NIN : 1][32][33];
  pdl_prologue();
  uint32_t bid = blockIdx.x;
  const uint32_t ti = bid % a.tilesI; bid /= a.tilesI;
  const uint32_t tj = bid % a.tilesJ; bid /= a.tilesJ;
  int64_t o_out = 0;
  int64_t o_in[NIN > 0 ? NIN : 1];
#pragma unroll
  for (int j = 0; j < NIN; ++j) o_in[j] = 0;
  for (int d = a.nbatch - 1; d >= 0; --d) {
    const uint32_t q = bid / a.bshape[d];
    const uint32_t r = bid - q * a.bshape[d];
    bid = q;
    o_out += (int64_t)r * a.bstride_out[d];
#pragma unroll
    for (int j = 0; j < NIN; ++j) o_in[j] += (int64_t)r * a.bstride_in[j][d];
  }
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int64_t i0 = (int64_t)ti * 32, j0 = (int64_t)tj * 32;
  // phase 1: staged inputs, threads run along J (their contiguous axis)
#pragma unroll
  for (int j = 0; j < NIN; ++j) {
    if (a.tr_mask & (1u << j)) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int ii = ty + k * 8;           // local I
        const int64_t gi = i0 + ii, gj = j0 + tx;
        if (gi < a.I && gj < a.J) tile[j][ii][tx] = a.in[j][o_in[j] + gi * a.sI_in[j] + gj * a.sJ_in[j]];
      }
    }
  }
  __syncthreads();
  // phase 2: threads run along I (output's contiguous axis)
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int jj = ty + k * 8;               // local J
    const int64_t gi = i0 + tx, gj = j0 + jj;
    if (gi < a.I && gj < a.J) {
      float x[NIN > 0 ? NIN : 1];
#pragma unroll
      for (int j = 0; j < NIN; ++j) {
        if (a.tr_mask & (1u << j)) x[j] = tile[j][tx][jj];
        else x[j] = a.in[j][o_in[j] + gi * a.sI_in[j] + gj * a.sJ_in[j]];
      }
      a.out[o_out + gi * a.sI_out + gj * a.sJ_out] = f(x);
    }
  }
}

// Vectorised variant: 64(J) x 64(I) tile, 128-bit global accesses on both sides, 16 elements per
// thread in flight.  Needs I % 4 == 0, J % 4 == 0 and 16-byte aligned rows for every operand.
template <int NIN, class F>
__global__ void __launch_bounds__(256) ew_tr4_kernel(const F f, const __grid_constant__ TrArgs<NIN> a) {
  __shared__ float tile[NIN > 0 ? NIN : 1][64][65];   // [i][j], pitch 65: conflict-light both ways
  pdl_prologue();
  uint32_t bid = blockIdx.x;
  const uint32_t ti = bid % a.tilesI; bid /= a.tilesI;
  const uint32_t tj = bid % a.tilesJ; bid /= a.tilesJ;
  int64_t o_out = 0;
  int64_t o_in[NIN > 0 ? NIN : 1];
#pragma unroll
  for (int j = 0; j < NIN; ++j) o_in[j] = 0;
  for (int d = a.nbatch - 1; d >= 0; --d) {
    const uint32_t q = bid / a.bshape[d];
    const uint32_t r = bid - q * a.bshape[d];
    bid = q;
    o_out += (int64_t)r * a.bstride_out[d];
#pragma unroll
    for (int j = 0; j < NIN; ++j) o_in[j] += (int64_t)r * a.bstride_in[j][d];
  }
  const int lx = (threadIdx.x & 15) * 4;   // 4-wide column group inside the tile
  const int ly = threadIdx.x >> 4;         // 0..15
  const int64_t i0 = (int64_t)ti * 64, j0 = (int64_t)tj * 64;
  // phase 1: staged inputs are contiguous along J: float4 loads along J, scattered to tile[i][j..j+3]
#pragma unroll
  for (int j = 0; j < NIN; ++j) {
    if (a.tr_mask & (1u << j)) {
      float4 t[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t gi = i0 + ly + 16 * k, gj = j0 + lx;
        t[k] = (gi < a.I && gj < a.J)
                   ? *reinterpret_cast<const float4*>(a.in[j] + o_in[j] + gi * a.sI_in[j] + gj)
                   : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float* row = &tile[j][ly + 16 * k][lx];
        row[0] = t[k].x; row[1] = t[k].y; row[2] = t[k].z; row[3] = t[k].w;
      }
    }
  }
  __syncthreads();
  // phase 2: float4 along I for the output and the non-staged inputs
  float4 direct[NIN > 0 ? NIN : 1][4];
#pragma unroll
  for (int j = 0; j < NIN; ++j) {
    if (!(a.tr_mask & (1u << j))) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t gj = j0 + ly + 16 * k, gi = i0 + lx;
        if (gi < a.I && gj < a.J) {
          if (a.sI_in[j] == 0) {
            const float sv = a.in[j][o_in[j] + gj * a.sJ_in[j]];
            direct[j][k] = make_float4(sv, sv, sv, sv);
          } else {
            direct[j][k] = *reinterpret_cast<const float4*>(a.in[j] + o_in[j] + gj * a.sJ_in[j] + gi);
          }
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int jj = ly + 16 * k;
    const int64_t gj = j0 + jj, gi = i0 + lx;
    if (gi < a.I && gj < a.J) {
      float r[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float x[NIN > 0 ? NIN : 1];
#pragma unroll
        for (int j = 0; j < NIN; ++j) {
          if (a.tr_mask & (1u << j)) x[j] = tile[j][lx + e][jj];
          else x[j] = (e == 0) ? direct[j][k].x : (e == 1) ? direct[j][k].y : (e == 2) ? direct[j][k].z : direct[j][k].w;
        }
        r[e] = f(x);
      }
      *reinterpret_cast<float4*>(a.out + o_out + gj * a.sJ_out + gi) = make_float4(r[0], r[1], r[2], r[3]);
    }
  }
}

// ---------------------------------------------------------------- dispatch
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int NIN, class F>
int launch_ew(const F& f, const wf2cu_view* out, const wf2cu_view* const* ins, const char* who) {
  WF2_REQUIRE_INIT();
  const wf2cu_view* views[1 + (NIN > 0 ? NIN : 1)];
  views[0] = out;
  for (int j = 0; j < NIN; ++j) views[1 + j] = ins[j];
  Plan p;
  int rc = build_plan(views, 1 + NIN, &p, who);
  if (rc) return rc;
  if (p.numel == 0) return 0;
  for (int d = 0; d < p.rank; ++d)
    if (p.stride[0][d] == 0 && p.shape[d] > 1)
      return set_error(WF2CU_ERR_INVALID, "%s: destination has a broadcast (stride-0) axis of extent %lld; use the reducing entry points",
                       who, (long long)p.shape[d]);
  State& s = S();
  float* optr = fptr(out);
  const float* iptr[NIN > 0 ? NIN : 1];
  for (int j = 0; j < NIN; ++j) iptr[j] = fptr(ins[j]);

  const int last = p.rank - 1;
  // ---- vec path?
  bool vec_ok = (p.rank <= 5) && (p.shape[last] % 4 == 0) && p.stride[0][last] == 1 && aligned16(optr) &&
                (p.numel / 4 < (int64_t(1) << 31));
  if (vec_ok) {
    for (int d = 0; d < last; ++d) if (p.stride[0][d] % 4 != 0) vec_ok = false;
    for (int j = 0; j < NIN && vec_ok; ++j) {
      const int64_t si = p.stride[1 + j][last];
      if (si == 1) {
        if (!aligned16(iptr[j])) vec_ok = false;
        for (int d = 0; d < last; ++d) if (p.stride[1 + j][d] % 4 != 0) vec_ok = false;
      } else if (si != 0) {
        vec_ok = false;
      }
    }
    if (p.rank > 1 && (p.numel / p.shape[last]) >= (int64_t(1) << 31)) vec_ok = false;
  }
  if (vec_ok) {
    VecArgs<NIN> a;
    memset(&a, 0, sizeof(a));
    a.out = optr;
    for (int j = 0; j < NIN; ++j) a.in[j] = iptr[j];
    a.outer_rank = p.rank - 1;
    a.inner4 = (uint32_t)(p.shape[last] / 4);
    for (int d = 0; d < last; ++d) {
      a.oshape[d] = (uint32_t)p.shape[d];
      a.ostride_out[d] = p.stride[0][d];
      for (int j = 0; j < NIN; ++j) a.ostride_in[j][d] = p.stride[1 + j][d];
    }
    for (int j = 0; j < NIN; ++j) if (p.stride[1 + j][last] == 0) a.inner_bcast |= (1u << j);
    a.n4 = p.numel / 4;
    const int64_t per_block = kThreads * kUnroll;
    const int64_t grid = (a.n4 + per_block - 1) / per_block;
    WF2_CUDA(launch_pdl(ew_vec_kernel<NIN, F>, dim3((unsigned)grid), dim3(kThreads), 0, s.stream, f, a));
    return post_launch(who);
  }
  // ---- transposed-operand path?
  if (NIN >= 1 && p.rank >= 2 && p.rank <= 5 && p.stride[0][last] == 1) {
    // find J: an axis != last where at least one input has stride 1 while its inner stride != 1
    int J = -1;
    for (int j = 0; j < NIN && J < 0; ++j)
      if (p.stride[1 + j][last] != 1 && p.stride[1 + j][last] != 0)
        for (int d = 0; d < last; ++d)
          if (p.stride[1 + j][d] == 1 && p.shape[d] >= 16) { J = d; break; }
    if (J >= 0 && p.shape[last] >= 16) {
      TrArgs<NIN> a;
      memset(&a, 0, sizeof(a));
      a.out = optr;
      for (int j = 0; j < NIN; ++j) a.in[j] = iptr[j];
      a.I = p.shape[last]; a.J = p.shape[J];
      a.sI_out = 1; a.sJ_out = p.stride[0][J];
      int nb = 0;
      int64_t nbatch_total = 1;
      for (int d = 0; d < last; ++d) {
        if (d == J) continue;
        a.bshape[nb] = (uint32_t)p.shape[d];
        a.bstride_out[nb] = p.stride[0][d];
        for (int j = 0; j < NIN; ++j) a.bstride_in[j][nb] = p.stride[1 + j][d];
        nbatch_total *= p.shape[d];
        ++nb;
      }
      a.nbatch = nb;
      for (int j = 0; j < NIN; ++j) {
        a.sI_in[j] = p.stride[1 + j][last];
        a.sJ_in[j] = p.stride[1 + j][J];
        if (a.sJ_in[j] == 1 && a.sI_in[j] != 1 && a.sI_in[j] != 0) a.tr_mask |= (1u << j);
      }
      // 128-bit variant: every extent/stride a multiple of 4 and 16-byte aligned bases
      bool v4 = (a.I % 4 == 0) && (a.J % 4 == 0) && aligned16(optr) && (a.sJ_out % 4 == 0) && a.I >= 64 && a.J >= 64;
      for (int b = 0; b < nb && v4; ++b) if (a.bstride_out[b] % 4) v4 = false;
      for (int j = 0; j < NIN && v4; ++j) {
        if (!aligned16(iptr[j])) v4 = false;
        for (int b = 0; b < nb; ++b) if (a.bstride_in[j][b] % 4) v4 = false;
        if (a.tr_mask & (1u << j)) { if (a.sI_in[j] % 4) v4 = false; }
        else if (a.sI_in[j] == 1) { if (a.sJ_in[j] % 4) v4 = false; }
        else if (a.sI_in[j] != 0) v4 = false;
      }
      const int tile = v4 ? 64 : 32;
      a.tilesI = (uint32_t)((a.I + tile - 1) / tile);
      a.tilesJ = (uint32_t)((a.J + tile - 1) / tile);
      const int64_t grid = (int64_t)a.tilesI * a.tilesJ * nbatch_total;
      if (nb <= 3 && grid < (int64_t(1) << 31)) {
        if (v4) WF2_CUDA(launch_pdl(ew_tr4_kernel<NIN, F>, dim3((unsigned)grid), dim3(256), 0, s.stream, f, a));
        else WF2_CUDA(launch_pdl(ew_tr_kernel<NIN, F>, dim3((unsigned)grid), dim3(256), 0, s.stream, f, a));
        return post_launch(who);
      }
    }
  }
  // ---- generic
  GenArgs<NIN> a;
  memset(&a, 0, sizeof(a));
  a.out = optr;
  for (int j = 0; j < NIN; ++j) a.in[j] = iptr[j];
  a.rank = p.rank;
  for (int d = 0; d < p.rank; ++d) {
    a.shape[d] = p.shape[d];
    a.stride_out[d] = p.stride[0][d];
    for (int j = 0; j < NIN; ++j) a.stride_in[j][d] = p.stride[1 + j][d];
  }
  a.numel = p.numel;
  int64_t grid = (p.numel + kThreads - 1) / kThreads;
  const int64_t cap = (int64_t)s.sm_count * 32;
  if (grid > cap) grid = cap;
  WF2_CUDA(launch_pdl(ew_gen_kernel<NIN, F>, dim3((unsigned)grid), dim3(kThreads), 0, s.stream, f, a));
  return post_launch(who);
}

// reducing binary (dst has stride-0 axes) lives in reduce.cu
int reduce_binary(int op, const wf2cu_view* dst, const wf2cu_view* src);
bool has_reduction_axes(const wf2cu_view* dst);


}  // namespace wf2

using namespace wf2;

extern "C" {

int wf2cu_ew_binary(int op, const wf2cu_view* dst, const wf2cu_view* src) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dst && src, "wf2cu_ew_binary: null view");
  if (has_reduction_axes(dst)) return reduce_binary(op, dst, src);
  const wf2cu_view* ins[2] = {dst, src};
  switch (op) {
    case WF2CU_ADD: return launch_ew<2>(FBinary<WF2CU_ADD>{}, dst, ins, "wf2cu_ew_binary(add)");
    case WF2CU_SUB: return launch_ew<2>(FBinary<WF2CU_SUB>{}, dst, ins, "wf2cu_ew_binary(sub)");
    case WF2CU_MUL: return launch_ew<2>(FBinary<WF2CU_MUL>{}, dst, ins, "wf2cu_ew_binary(mul)");
    case WF2CU_DIV: return launch_ew<2>(FBinary<WF2CU_DIV>{}, dst, ins, "wf2cu_ew_binary(div)");
    case WF2CU_COPY: { const wf2cu_view* one[1] = {src}; return launch_ew<1>(FCopy{}, dst, one, "wf2cu_ew_binary(copy)"); }
    default: return set_error(WF2CU_ERR_INVALID, "wf2cu_ew_binary: unknown op %d", op);
  }
}

int wf2cu_ew_binary3(int op, const wf2cu_view* out, const wf2cu_view* a, const wf2cu_view* b) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(out && a && b, "wf2cu_ew_binary3: null view");
  const wf2cu_view* ins[2] = {a, b};
  switch (op) {
    case WF2CU_ADD: return launch_ew<2>(FBinary<WF2CU_ADD>{}, out, ins, "wf2cu_ew_binary3(add)");
    case WF2CU_SUB: return launch_ew<2>(FBinary<WF2CU_SUB>{}, out, ins, "wf2cu_ew_binary3(sub)");
    case WF2CU_MUL: return launch_ew<2>(FBinary<WF2CU_MUL>{}, out, ins, "wf2cu_ew_binary3(mul)");
    case WF2CU_DIV: return launch_ew<2>(FBinary<WF2CU_DIV>{}, out, ins, "wf2cu_ew_binary3(div)");
    case WF2CU_COPY: { const wf2cu_view* one[1] = {b}; return launch_ew<1>(FCopy{}, out, one, "wf2cu_ew_binary3(copy)"); }
    default: return set_error(WF2CU_ERR_INVALID, "wf2cu_ew_binary3: unknown op %d", op);
  }
}

int wf2cu_ew_scalar(int op, const wf2cu_view* dst, float scalar) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dst, "wf2cu_ew_scalar: null view");
  const wf2cu_view* ins[1] = {dst};
  switch (op) {
    case WF2CU_ADD: return launch_ew<1>(FScalar<WF2CU_ADD>{scalar}, dst, ins, "wf2cu_ew_scalar(add)");
    case WF2CU_SUB: return launch_ew<1>(FScalar<WF2CU_SUB>{scalar}, dst, ins, "wf2cu_ew_scalar(sub)");
    case WF2CU_MUL: return launch_ew<1>(FScalar<WF2CU_MUL>{scalar}, dst, ins, "wf2cu_ew_scalar(mul)");
    case WF2CU_DIV: return launch_ew<1>(FScalar<WF2CU_DIV>{scalar}, dst, ins, "wf2cu_ew_scalar(div)");
    case WF2CU_COPY: return launch_ew<0>(FFill{scalar}, dst, nullptr, "wf2cu_ew_scalar(copy)");
    default: return set_error(WF2CU_ERR_INVALID, "wf2cu_ew_scalar: unknown op %d", op);
  }
}

int wf2cu_fill(const wf2cu_view* dst, float value) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dst, "wf2cu_fill: null view");
  return launch_ew<0>(FFill{value}, dst, nullptr, "wf2cu_fill");
}

int wf2cu_ew_unary(int fn, const wf2cu_view* src, const wf2cu_view* dst) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dst && src, "wf2cu_ew_unary: null view");
  const wf2cu_view* ins[1] = {src};
#define WF2_CASE(FN) case FN: return launch_ew<1>(FUnary<FN>{}, dst, ins, "wf2cu_ew_unary(" #FN ")");
  switch (fn) {
    WF2_CASE(WF2CU_SIN) WF2_CASE(WF2CU_COS) WF2_CASE(WF2CU_TAN) WF2_CASE(WF2CU_ASIN) WF2_CASE(WF2CU_ACOS)
    WF2_CASE(WF2CU_ATAN) WF2_CASE(WF2CU_SINH) WF2_CASE(WF2CU_COSH) WF2_CASE(WF2CU_TANH) WF2_CASE(WF2CU_ABS)
    WF2_CASE(WF2CU_EXP) WF2_CASE(WF2CU_LOGE) WF2_CASE(WF2CU_LOG2) WF2_CASE(WF2CU_LOG10) WF2_CASE(WF2CU_LOG1P)
    WF2_CASE(WF2CU_SQRT) WF2_CASE(WF2CU_CBRT) WF2_CASE(WF2CU_SIGN) WF2_CASE(WF2CU_SQUARE) WF2_CASE(WF2CU_NEG)
    WF2_CASE(WF2CU_RECIP)
    default: return set_error(WF2CU_ERR_INVALID, "wf2cu_ew_unary: unknown function %d", fn);
  }
#undef WF2_CASE
}

int wf2cu_pow(const wf2cu_view* src, const wf2cu_view* dst, float n) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dst && src, "wf2cu_pow: null view");
  const wf2cu_view* ins[1] = {src};
  return launch_ew<1>(FPow{n}, dst, ins, "wf2cu_pow");
}

int wf2cu_cmp(int op, const wf2cu_view* x, const wf2cu_view* y, float scal, const wf2cu_view* dst,
              float then_value, float else_value) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dst && x, "wf2cu_cmp: null view");
  WF2_CHECK(op >= WF2CU_EQ && op <= WF2CU_GE, "wf2cu_cmp: unknown compare %d", op);
  if (y) {
    const wf2cu_view* ins[2] = {x, y};
    return launch_ew<2>(FCmp2{op, then_value, else_value}, dst, ins, "wf2cu_cmp");
  }
  const wf2cu_view* ins[1] = {x};
  return launch_ew<1>(FCmp1{op, scal, then_value, else_value}, dst, ins, "wf2cu_cmp(scal)");
}

}  // extern "C"
