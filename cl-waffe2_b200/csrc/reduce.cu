// reduce.cu — `!sum` / `!mean` and every other "AddNode whose destination has stride 0".
//
// In the reference a reduction is not a node: `!sum` zero-fills `out` with `A*=scal out 0`, views
// it with `(:broadcast n)` on the reduced axes and runs `A+=B out* x`
// (`source/base-impl/reduction.lisp:60-79`); the C kernel then sees incx = 0, skips its SIMD body
// and accumulates `x[0] = x[0] + y[i]` sequentially (`cl-waffe2-simd.c:159,205-212`).  A parallel
// element-wise kernel on such a destination would race, so wf2cu_ew_binary classifies by the
// destination strides and lands here.  Summation is a fixed-shape tree (warp shuffle -> block ->
// optional second stage), i.e. deterministic, and differs from the reference's sequential order
// by O(log n) roundings instead of O(n) (SURVEY §7 hard part d).
//
// Kernels (all HBM-bound, one pass over x):
//   rows : kept axes outermost, ONE reduced axis innermost & contiguous  (sum :axis 1, :axis t)
//   cols : [batch,] ONE reduced axis, then kept contiguous axis          (sum :axis 0, 4-D :axis 1)
//   gen  : any other mix of kept/reduced axes (warp per output element)
#include "common.cuh"
#include "math.cuh"
#include <cooperative_groups.h>
#include <stdlib.h>
namespace cg = cooperative_groups;

namespace wf2 {

struct MSum {
  __device__ __forceinline__ static float id() { return 0.f; }
  __device__ __forceinline__ static float op(float a, float b) { return a + b; }   // no multiply nearby: nothing to contract
};
struct MProd {
  __device__ __forceinline__ static float id() { return 1.f; }
  __device__ __forceinline__ static float op(float a, float b) { return a * b; }
};

// how the reduced total t meets the old destination value d
enum { FIN_ADD = 0, FIN_SUB, FIN_MUL, FIN_DIV, FIN_SET };
struct Fin {
  int mode;
  float divisor;  // applied last when != 0
  __device__ __forceinline__ float operator()(float d, float t) const {
    float r;
    switch (mode) {
      case FIN_ADD: r = __fadd_rn(d, t); break;
      case FIN_SUB: r = __fsub_rn(d, t); break;
      case FIN_MUL: r = __fmul_rn(d, t); break;
      case FIN_DIV: r = __fdiv_rn(d, t); break;
      default: r = t;
    }
    if (divisor != 0.f) r = __fdiv_rn(r, divisor);
    return r;
  }
  // FIN_SET never needs the old destination value.  This matters: a dependent read of dst at the END
  // of a CTA (one L2/HBM round trip queued behind the streaming loads) kept every CTA resident ~1 us
  // longer and cost 6 us of 17 us on the (4096 4096) row sum (gpurun_out/redbench5.log, v4 vs old).
  __device__ __forceinline__ bool reads_dst() const { return mode != FIN_SET; }
  // for writers that cannot prefetch the old value (ticket winners)
  __device__ __forceinline__ void store(float* p, float t) const {
    const float d = reads_dst() ? *p : 0.f;
    *p = (*this)(d, t);
  }
};

// 128-bit global load the compiler may not sink below later code: volatile asm statements keep their
// program order, so a run of these is issued back to back (the loads-in-flight count is then what
// the source says, not what ptxas' register-pressure heuristic picks).
__device__ __forceinline__ float4 ld_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

template <class M>
__device__ __forceinline__ float warp_reduce(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = M::op(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

template <class M, int THREADS>
__device__ __forceinline__ float block_reduce(float v, float* smem /* THREADS/32 floats */) {
  v = warp_reduce<M>(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) smem[w] = v;
  __syncthreads();
  if (w == 0) {
    float t = (lane < THREADS / 32) ? smem[lane] : M::id();
    t = warp_reduce<M>(t);
    if (lane == 0) smem[0] = t;
  }
  __syncthreads();
  const float r = smem[0];
  __syncthreads();
  return r;
}

// ------------------------------------------------------------------ rows
struct RowArgs {
  const float* src;
  float* dst;        // final destination (kept-axes strided)
  float* partial;    // [rows * chunks] when chunks > 1
  int nkept;         // 0..4 kept axes
  uint32_t kshape[4];
  int64_t kstride_src[4], kstride_dst[4];
  int64_t cols;      // reduced extent, src stride 1
  int64_t chunk_len; // columns per chunk (multiple of 4)
  uint32_t chunks;
  int vec;           // 1: float4 loads legal for every row
  unsigned int* counters;  // per-row tickets for the fused second stage (chunks > 1)
};

template <class M, int THREADS>
__global__ void __launch_bounds__(THREADS) reduce_rows_kernel(const __grid_constant__ RowArgs a, const Fin fin) {
  __shared__ float smem[THREADS / 32];
  const uint32_t chunk = blockIdx.x % a.chunks;
  uint32_t row = blockIdx.x / a.chunks;
  const uint32_t row_id = row;
  int64_t so = 0, dofs = 0;
  for (int d = a.nkept - 1; d >= 0; --d) {
    const uint32_t q = row / a.kshape[d];
    const uint32_t r = row - q * a.kshape[d];
    row = q;
    so += (int64_t)r * a.kstride_src[d];
    dofs += (int64_t)r * a.kstride_dst[d];
  }
  const int64_t c0 = (int64_t)chunk * a.chunk_len;
  int64_t c1 = c0 + a.chunk_len;
  if (c1 > a.cols) c1 = a.cols;
  const float* p = a.src + so;
  float d_old = 0.f;   // accumulate modes: fetch the old destination value now, under the main loads
  if (a.chunks == 1 && threadIdx.x == 0 && fin.reads_dst()) d_old = a.dst[dofs];
  float acc8[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc8[k] = M::id();
  if (a.vec) {
    const int64_t v0 = c0 >> 2, v1 = c1 >> 2;  // c0 multiple of 4; tail handled below
    const float4* p4 = reinterpret_cast<const float4*>(p);
    int64_t i = v0 + threadIdx.x;
    // unroll 1: when nvcc unrolls this loop itself, ptxas serialises the loads through one register quad
#pragma unroll 1
    for (; i + 7 * THREADS < v1; i += 8 * THREADS) {
      float4 x[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) x[k] = p4[i + k * THREADS];
#pragma unroll
      for (int k = 0; k < 8; ++k) acc8[k] = M::op(acc8[k], M::op(M::op(x[k].x, x[k].y), M::op(x[k].z, x[k].w)));
    }
#pragma unroll 1
    for (; i + 3 * THREADS < v1; i += 4 * THREADS) {
      float4 x[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) x[k] = p4[i + k * THREADS];
#pragma unroll
      for (int k = 0; k < 4; ++k) acc8[k] = M::op(acc8[k], M::op(M::op(x[k].x, x[k].y), M::op(x[k].z, x[k].w)));
    }
    for (; i < v1; i += THREADS) {
      const float4 x0 = p4[i];
      acc8[0] = M::op(acc8[0], M::op(M::op(x0.x, x0.y), M::op(x0.z, x0.w)));
    }
    for (int64_t c = (v1 << 2) + threadIdx.x; c < c1; c += THREADS) acc8[1] = M::op(acc8[1], p[c]);
  } else {
    for (int64_t c = c0 + threadIdx.x; c < c1; c += THREADS) acc8[0] = M::op(acc8[0], p[c]);
  }
  const float acc0 = M::op(acc8[0], acc8[4]), acc1 = M::op(acc8[1], acc8[5]);
  const float acc2 = M::op(acc8[2], acc8[6]), acc3 = M::op(acc8[3], acc8[7]);
  float acc = M::op(M::op(acc0, acc1), M::op(acc2, acc3));
  acc = block_reduce<M, THREADS>(acc, smem);
  if (a.chunks == 1) {
    if (threadIdx.x == 0) a.dst[dofs] = fin(d_old, acc);
    return;
  }
  // chunks > 1: publish the partial; the LAST CTA of this row folds all partials in a fixed order
  // (deterministic) and resets the ticket, so no second launch is needed.
  __shared__ bool is_last;
  if (threadIdx.x == 0) {
    float* slot = a.partial + (int64_t)row_id * a.chunks + chunk;
    unsigned int t;
    asm volatile("st.relaxed.gpu.global.f32 [%0], %1;" ::"l"(slot), "f"(acc) : "memory");
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(a.counters + row_id) : "memory");
    is_last = (t == a.chunks - 1);
  }
  __syncthreads();
  if (!is_last) return;
  const float* pp = a.partial + (int64_t)row_id * a.chunks;
  float v = M::id();
  for (uint32_t c = threadIdx.x; c < a.chunks; c += THREADS) v = M::op(v, __ldcg(pp + c));
  v = block_reduce<M, THREADS>(v, smem);
  if (threadIdx.x == 0) {
    fin.store(a.dst + dofs, v);
    a.counters[row_id] = 0;
  }
}

// several short rows per CTA: one warp per row
template <class M>
__global__ void __launch_bounds__(256) reduce_rows_warp_kernel(const __grid_constant__ RowArgs a, const Fin fin, uint32_t nrows) {
  const uint32_t row_id = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row_id >= nrows) return;
  const int lane = threadIdx.x & 31;
  uint32_t row = row_id;
  int64_t so = 0, dofs = 0;
  for (int d = a.nkept - 1; d >= 0; --d) {
    const uint32_t q = row / a.kshape[d];
    const uint32_t r = row - q * a.kshape[d];
    row = q;
    so += (int64_t)r * a.kstride_src[d];
    dofs += (int64_t)r * a.kstride_dst[d];
  }
  const float* p = a.src + so;
  float acc = M::id();
  float d_old = 0.f;
  if (lane == 0 && fin.reads_dst()) d_old = a.dst[dofs];
  if (a.vec) {
    const float4* p4 = reinterpret_cast<const float4*>(p);
    const int64_t v1 = a.cols >> 2;
    float b1 = M::id(), b2 = M::id(), b3 = M::id();
    int64_t i = lane;
    for (; i + 96 < v1; i += 128) {
      const float4 x0 = p4[i], x1 = p4[i + 32], x2 = p4[i + 64], x3 = p4[i + 96];
      acc = M::op(acc, M::op(M::op(x0.x, x0.y), M::op(x0.z, x0.w)));
      b1 = M::op(b1, M::op(M::op(x1.x, x1.y), M::op(x1.z, x1.w)));
      b2 = M::op(b2, M::op(M::op(x2.x, x2.y), M::op(x2.z, x2.w)));
      b3 = M::op(b3, M::op(M::op(x3.x, x3.y), M::op(x3.z, x3.w)));
    }
    for (; i < v1; i += 32) {
      const float4 x = p4[i];
      acc = M::op(acc, M::op(M::op(x.x, x.y), M::op(x.z, x.w)));
    }
    acc = M::op(M::op(acc, b1), M::op(b2, b3));
    for (int64_t c = (v1 << 2) + lane; c < a.cols; c += 32) acc = M::op(acc, p[c]);
  } else {
    for (int64_t c = lane; c < a.cols; c += 32) acc = M::op(acc, p[c]);
  }
  acc = warp_reduce<M>(acc);
  if (lane == 0) a.dst[dofs] = fin(d_old, acc);
}

// Tile kernel for rows that fit a register tile: a CTA of 256 threads takes RPC rows, TPR = 256/RPC
// threads per row, and every thread issues ALL of its (up to 8) 128-bit loads before the first add,
// so a resident SM has ~256 KB of loads in flight.  Needs a.vec and cols/4 <= TPR * 8.
template <class M, int RPC>
__global__ void __launch_bounds__(256) reduce_rows_tile_kernel(const __grid_constant__ RowArgs a, const Fin fin, uint32_t nrows) {
  constexpr int TPR = 256 / RPC;
  constexpr int WPR = TPR / 32;                 // warps per row
  __shared__ float smem[8];
  pdl_prologue();
  const int sub = threadIdx.x / TPR;            // row slot inside the CTA
  const int t = threadIdx.x % TPR;
  const uint32_t row_id = blockIdx.x * RPC + sub;
  const bool live = row_id < nrows;
  int64_t so = 0, dofs = 0;
  uint32_t row = live ? row_id : 0;
  for (int d = a.nkept - 1; d >= 0; --d) {
    const uint32_t q = row / a.kshape[d];
    const uint32_t r = row - q * a.kshape[d];
    row = q;
    so += (int64_t)r * a.kstride_src[d];
    dofs += (int64_t)r * a.kstride_dst[d];
  }
  const float4* p4 = reinterpret_cast<const float4*>(a.src + so);
  const int64_t v1 = a.cols >> 2;
  float d_old = 0.f;
  if (live && t == 0 && fin.reads_dst()) d_old = a.dst[dofs];
  // Loads are UNCONDITIONAL (index clamped into the row; dead slots read row 0) and the identity is
  // selected afterwards: with predicated loads ptxas issued them in two batches of 4 (32 registers),
  // which cost 6 us of the 17 us this kernel took on (4096 4096) (profiles/r01_launches_summary.md).
  const int vlast = (int)v1 - 1;
  float4 x[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) x[k] = ld_f4(p4 + min(t + k * TPR, vlast));
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float sk = M::op(M::op(x[k].x, x[k].y), M::op(x[k].z, x[k].w));
    acc[k] = (live && t + k * TPR <= vlast) ? sk : M::id();
  }
  float v = M::op(M::op(M::op(acc[0], acc[4]), M::op(acc[1], acc[5])), M::op(M::op(acc[2], acc[6]), M::op(acc[3], acc[7])));
  if (live) {
    const float* p = a.src + so;
    for (int64_t c = (v1 << 2) + t; c < a.cols; c += TPR) v = M::op(v, p[c]);
  }
  v = warp_reduce<M>(v);
  if (WPR == 1) {
    if (live && (threadIdx.x & 31) == 0) a.dst[dofs] = fin(d_old, v);
    return;
  }
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) smem[w] = v;
  __syncthreads();
  if (t == 0 && live) {
    float r = smem[sub * WPR];
#pragma unroll
    for (int k = 1; k < WPR; ++k) r = M::op(r, smem[sub * WPR + k]);
    a.dst[dofs] = fin(d_old, r);
  }
}

// ------------------------------------------------------------------ flat (sum :axis t of a contiguous tensor)
// One long contiguous run, no kept axes.  ptxas keeps only two 128-bit register loads in flight per
// thread in the looped row kernel above however the source is written (the SASS re-uses two register
// quads), so here the loads do not go through registers at all: a persistent CTA (2 per SM) streams
// 16 KiB chunks into a 3-stage shared-memory ring with the bulk async-copy engine
// (cp.async.bulk.shared.global + mbarrier complete_tx), i.e. 48 KiB in flight per CTA by construction.
// Partials meet through a ticket (st.relaxed + atom.acq_rel); the last CTA folds them in a fixed order.
constexpr int kFlatThreads = 256, kFlatStages = 4, kFlatChunk = 16384;   // bytes per stage
struct FlatArgs {
  const float* src;     // 16-byte aligned
  float* dst;
  float* partial;       // [gridDim.x]
  unsigned int* counter;
  int64_t n;            // elements
  int stages;           // ring depth actually used (<= kFlatStages; the launch passes stages * kFlatChunk bytes)
  int late_trigger;     // 1: let the dependent grid in only when this CTA has finished streaming
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <class M>
__global__ void __launch_bounds__(kFlatThreads) reduce_flat_bulk_kernel(const __grid_constant__ FlatArgs a, const Fin fin) {
  extern __shared__ __align__(128) uint8_t ring[];
  __shared__ uint64_t bar[kFlatStages];
  __shared__ float red[kFlatThreads / 32];
  __shared__ bool is_last;
  constexpr int T = kFlatThreads;
  const int S = a.stages;
  const int64_t nchunks = (a.n * 4) / kFlatChunk;
  const char* base = reinterpret_cast<const char*>(a.src);
  // shared-memory set-up first: under programmatic dependent launch it overlaps the predecessor's tail
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(&bar[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (!a.late_trigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (threadIdx.x == 0) {
    int64_t c = blockIdx.x;
    for (int s = 0; s < S && c < nchunks; ++s, c += gridDim.x) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&bar[s])), "r"(kFlatChunk) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_addr(ring + s * kFlatChunk)), "l"(base + c * kFlatChunk), "r"(kFlatChunk), "r"(smem_addr(&bar[s])) : "memory");
    }
  }
  float acc0 = M::id(), acc1 = M::id(), acc2 = M::id(), acc3 = M::id();
  int s = 0;
  uint32_t ph = 0;
  for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\tselp.u32 %0,1,0,q;\n\t}"
                   : "=r"(ok) : "r"(smem_addr(&bar[s])), "r"(ph) : "memory");
    const float4* t = reinterpret_cast<const float4*>(ring + s * kFlatChunk);
    // 1024 float4 per chunk, 4 per thread, conflict-free (consecutive lanes, consecutive 16 B)
    const float4 x0 = t[threadIdx.x], x1 = t[threadIdx.x + T], x2 = t[threadIdx.x + 2 * T], x3 = t[threadIdx.x + 3 * T];
    acc0 = M::op(acc0, M::op(M::op(x0.x, x0.y), M::op(x0.z, x0.w)));
    acc1 = M::op(acc1, M::op(M::op(x1.x, x1.y), M::op(x1.z, x1.w)));
    acc2 = M::op(acc2, M::op(M::op(x2.x, x2.y), M::op(x2.z, x2.w)));
    acc3 = M::op(acc3, M::op(M::op(x3.x, x3.y), M::op(x3.z, x3.w)));
    __syncthreads();   // every thread is done with stage s
    if (threadIdx.x == 0) {
      const int64_t nc = c + (int64_t)S * gridDim.x;
      if (nc < nchunks) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(&bar[s])), "r"(kFlatChunk) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_addr(ring + s * kFlatChunk)), "l"(base + nc * kFlatChunk), "r"(kFlatChunk), "r"(smem_addr(&bar[s])) : "memory");
      }
    }
    if (++s == S) { s = 0; ph ^= 1; }
  }
  if (a.late_trigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // ragged tail (< one chunk): CTA 0, plain loads
  if (blockIdx.x == 0)
    for (int64_t i = nchunks * (kFlatChunk / 4) + threadIdx.x; i < a.n; i += T) acc0 = M::op(acc0, a.src[i]);
  float v = M::op(M::op(acc0, acc1), M::op(acc2, acc3));
  v = block_reduce<M, T>(v, red);
  if (threadIdx.x == 0) {
    unsigned int t;
    asm volatile("st.relaxed.gpu.global.f32 [%0], %1;" ::"l"(a.partial + blockIdx.x), "f"(v) : "memory");
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(a.counter) : "memory");
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  float w = M::id();
  for (uint32_t c = threadIdx.x; c < gridDim.x; c += T) w = M::op(w, __ldcg(a.partial + c));
  w = block_reduce<M, T>(w, red);
  if (threadIdx.x == 0) {
    fin.store(a.dst, w);
    *a.counter = 0;
  }
}

// ------------------------------------------------------------------ cols
struct ColArgs {
  const float* src;
  float* dst;
  float* partial;          // [batch][splits][K] when splits > 1
  int64_t B, R, K;         // batch, reduced, kept (contiguous)
  int64_t sB_src, sR_src;  // src strides (kept stride is 1)
  int64_t sB_dst;          // dst batch stride (kept stride is 1)
  uint32_t splits;         // CTAs along R
  int64_t rows_per_split;
  int vec;
  unsigned int* counters;  // per (batch, column-block) tickets: the last split folds the partials
};

// block = 32 (x: columns, float4 each when vec) x 8 (y: rows)
template <class M>
__global__ void __launch_bounds__(256) reduce_cols_kernel(const __grid_constant__ ColArgs a, const Fin fin) {
  __shared__ float4 sm[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t b = blockIdx.z;
  const uint32_t split = blockIdx.y;
  const int64_t r0 = (int64_t)split * a.rows_per_split;
  int64_t r1 = r0 + a.rows_per_split;
  if (r1 > a.R) r1 = a.R;
  const float* base = a.src + b * a.sB_src;
  float4 acc = make_float4(M::id(), M::id(), M::id(), M::id());
  float4 acc2 = acc;
  if (a.vec) {
    const int64_t col = ((int64_t)blockIdx.x * 32 + tx) * 4;
    if (col < a.K) {
      int64_t r = r0 + ty;
      float4 acc3 = acc2, acc4 = acc2;
      for (; r + 24 < r1; r += 32) {   // 4 independent 128-bit loads in flight
        const float4 x = *reinterpret_cast<const float4*>(base + r * a.sR_src + col);
        const float4 y = *reinterpret_cast<const float4*>(base + (r + 8) * a.sR_src + col);
        const float4 z = *reinterpret_cast<const float4*>(base + (r + 16) * a.sR_src + col);
        const float4 w = *reinterpret_cast<const float4*>(base + (r + 24) * a.sR_src + col);
        asm volatile("" ::: "memory");
        acc.x = M::op(acc.x, x.x); acc.y = M::op(acc.y, x.y); acc.z = M::op(acc.z, x.z); acc.w = M::op(acc.w, x.w);
        acc2.x = M::op(acc2.x, y.x); acc2.y = M::op(acc2.y, y.y); acc2.z = M::op(acc2.z, y.z); acc2.w = M::op(acc2.w, y.w);
        acc3.x = M::op(acc3.x, z.x); acc3.y = M::op(acc3.y, z.y); acc3.z = M::op(acc3.z, z.z); acc3.w = M::op(acc3.w, z.w);
        acc4.x = M::op(acc4.x, w.x); acc4.y = M::op(acc4.y, w.y); acc4.z = M::op(acc4.z, w.z); acc4.w = M::op(acc4.w, w.w);
      }
      for (; r < r1; r += 8) {
        const float4 x = *reinterpret_cast<const float4*>(base + r * a.sR_src + col);
        acc.x = M::op(acc.x, x.x); acc.y = M::op(acc.y, x.y); acc.z = M::op(acc.z, x.z); acc.w = M::op(acc.w, x.w);
      }
      acc2.x = M::op(acc2.x, M::op(acc3.x, acc4.x)); acc2.y = M::op(acc2.y, M::op(acc3.y, acc4.y));
      acc2.z = M::op(acc2.z, M::op(acc3.z, acc4.z)); acc2.w = M::op(acc2.w, M::op(acc3.w, acc4.w));
    }
    acc.x = M::op(acc.x, acc2.x); acc.y = M::op(acc.y, acc2.y); acc.z = M::op(acc.z, acc2.z); acc.w = M::op(acc.w, acc2.w);
  } else {
    // scalar columns: each thread owns column blockIdx.x*128 + tx + 32*k, k = 0..3 (kept in .x..w)
    const int64_t cb = (int64_t)blockIdx.x * 128 + tx;
    for (int64_t r = r0 + ty; r < r1; r += 8) {
      const float* rowp = base + r * a.sR_src;
      if (cb < a.K) acc.x = M::op(acc.x, rowp[cb]);
      if (cb + 32 < a.K) acc.y = M::op(acc.y, rowp[cb + 32]);
      if (cb + 64 < a.K) acc.z = M::op(acc.z, rowp[cb + 64]);
      if (cb + 96 < a.K) acc.w = M::op(acc.w, rowp[cb + 96]);
    }
  }
  sm[ty][tx] = acc;
  __syncthreads();
  const bool final_stage = (a.splits == 1);
  if (ty == 0) {
#pragma unroll
    for (int k = 1; k < 8; ++k) {
      const float4 o = sm[k][tx];
      acc.x = M::op(acc.x, o.x); acc.y = M::op(acc.y, o.y); acc.z = M::op(acc.z, o.z); acc.w = M::op(acc.w, o.w);
    }
    float* outp = final_stage ? a.dst + b * a.sB_dst : a.partial + (b * a.splits + split) * a.K;
    if (a.vec) {
      const int64_t col = ((int64_t)blockIdx.x * 32 + tx) * 4;
      if (col < a.K) {
        if (final_stage) {
          float4 d = fin.reads_dst() ? *reinterpret_cast<float4*>(outp + col) : make_float4(0.f, 0.f, 0.f, 0.f);
          d.x = fin(d.x, acc.x); d.y = fin(d.y, acc.y); d.z = fin(d.z, acc.z); d.w = fin(d.w, acc.w);
          *reinterpret_cast<float4*>(outp + col) = d;
        } else {
          *reinterpret_cast<float4*>(outp + col) = acc;
        }
      }
    } else {
      const int64_t cb = (int64_t)blockIdx.x * 128 + tx;
      const float vals[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int64_t c = cb + 32 * k;
        if (c < a.K) { if (final_stage) fin.store(outp + c, vals[k]); else outp[c] = vals[k]; }
      }
    }
  }
  if (final_stage || a.counters == nullptr) return;
  // fused second stage: the last split to arrive for this (batch, column block) folds all partials
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int t = atomicAdd(&a.counters[b * gridDim.x + blockIdx.x], 1u);
    is_last = (t == a.splits - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  {
    // columns of this block: 128; thread (tx, ty): column group tx (4 columns), splits ty, ty+8, ...
    const float* pb = a.partial + b * a.splits * a.K;
    float4 t4 = make_float4(M::id(), M::id(), M::id(), M::id());
    const int64_t colv = ((int64_t)blockIdx.x * 32 + tx) * 4;
    for (uint32_t sp = ty; sp < a.splits; sp += 8) {
      const float* q = pb + (int64_t)sp * a.K;
      if (a.vec) {
        if (colv < a.K) {
          const float4 o = __ldcg(reinterpret_cast<const float4*>(q + colv));
          t4.x = M::op(t4.x, o.x); t4.y = M::op(t4.y, o.y); t4.z = M::op(t4.z, o.z); t4.w = M::op(t4.w, o.w);
        }
      } else {
        const int64_t cb = (int64_t)blockIdx.x * 128 + tx;
        if (cb < a.K) t4.x = M::op(t4.x, __ldcg(q + cb));
        if (cb + 32 < a.K) t4.y = M::op(t4.y, __ldcg(q + cb + 32));
        if (cb + 64 < a.K) t4.z = M::op(t4.z, __ldcg(q + cb + 64));
        if (cb + 96 < a.K) t4.w = M::op(t4.w, __ldcg(q + cb + 96));
      }
    }
    sm[ty][tx] = t4;
    __syncthreads();
    if (ty == 0) {
#pragma unroll
      for (int k = 1; k < 8; ++k) {
        const float4 o = sm[k][tx];
        t4.x = M::op(t4.x, o.x); t4.y = M::op(t4.y, o.y); t4.z = M::op(t4.z, o.z); t4.w = M::op(t4.w, o.w);
      }
      float* outp = a.dst + b * a.sB_dst;
      if (a.vec) {
        if (colv < a.K) {
          float4 d = fin.reads_dst() ? *reinterpret_cast<float4*>(outp + colv) : make_float4(0.f, 0.f, 0.f, 0.f);
          d.x = fin(d.x, t4.x); d.y = fin(d.y, t4.y); d.z = fin(d.z, t4.z); d.w = fin(d.w, t4.w);
          *reinterpret_cast<float4*>(outp + colv) = d;
        }
      } else {
        const int64_t cb = (int64_t)blockIdx.x * 128 + tx;
        const float vals[4] = {t4.x, t4.y, t4.z, t4.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int64_t c = cb + 32 * k;
          if (c < a.K) fin.store(outp + c, vals[k]);
        }
      }
      if (tx == 0) a.counters[b * gridDim.x + blockIdx.x] = 0;
    }
  }
}

// Narrow variant: K <= 64 kept columns, rows dense (sR_src == K), many rows — bias gradients of layers with
// few outputs (Conv2D's (N*Ho*Wo, C_out) gradient: 67 600 x 4 on MNIST-CNN) — where a column-block grid
// would give one thread column per CTA.  The tensor is read as ONE flat contiguous run: CTA g owns rows
// [g*rps, ...), thread t < L*K (L = 256 / K) owns column t % K of rows t / K, t / K + L, ...; consecutive
// threads read consecutive floats.  Column totals meet in shared memory, CTA partials through the ticket,
// both in a fixed order.  52 us -> a few us on the shape above.
template <class M>
__global__ void __launch_bounds__(256) reduce_cols_narrow_kernel(const __grid_constant__ ColArgs a, const Fin fin) {
  __shared__ float sm[256];
  __shared__ bool is_last;
  const int K = (int)a.K;
  const int L = 256 / K;                 // rows per step
  const int t = threadIdx.x;
  const bool active = t < L * K;
  const int col = t % K, lr = t / K;
  const int64_t r0 = (int64_t)blockIdx.x * a.rows_per_split;
  int64_t r1 = r0 + a.rows_per_split;
  if (r1 > a.R) r1 = a.R;
  float acc0 = M::id(), acc1 = M::id(), acc2 = M::id(), acc3 = M::id();
  if (active) {
    const float* p = a.src + col;
    int64_t r = r0 + lr;
    for (; r + 3 * L < r1; r += 4 * L) {
      const float x0 = p[r * K], x1 = p[(r + L) * K], x2 = p[(r + 2 * L) * K], x3 = p[(r + 3 * L) * K];
      acc0 = M::op(acc0, x0); acc1 = M::op(acc1, x1); acc2 = M::op(acc2, x2); acc3 = M::op(acc3, x3);
    }
    for (; r < r1; r += L) acc0 = M::op(acc0, p[r * K]);
  }
  sm[t] = M::op(M::op(acc0, acc1), M::op(acc2, acc3));
  __syncthreads();
  float tot = M::id();
  if (t < K) {
    for (int k = 0; k < L; ++k) tot = M::op(tot, sm[k * K + t]);
    if (a.splits == 1) { fin.store(a.dst + t, tot); }
    else a.partial[(int64_t)blockIdx.x * K + t] = tot;
  }
  if (a.splits == 1) return;
  __syncthreads();
  if (t == 0) {
    __threadfence();
    is_last = (atomicAdd(&a.counters[0], 1u) == a.splits - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // the last CTA folds the partials: thread t owns column t % K of splits t / K, t / K + L, ...
  float f = M::id();
  if (active)
    for (uint32_t sp = lr; sp < a.splits; sp += L) f = M::op(f, __ldcg(a.partial + (int64_t)sp * K + col));
  sm[t] = f;
  __syncthreads();
  if (t < K) {
    float r = M::id();
    for (int k = 0; k < L; ++k) r = M::op(r, sm[k * K + t]);
    fin.store(a.dst + t, r);
  }
  if (t == 0) a.counters[0] = 0;
}

// Cluster variant (vec only): 8 CTAs along the reduced axis form a thread-block cluster, each sums its
// eighth of the rows for a 64-column strip (16 lanes x float4, 16 rows per step, 4 loads in flight), and
// the 8 partial strips meet in DISTRIBUTED SHARED MEMORY: rank 0 reads its peers' strips with
// ld.shared::cluster and writes the result.  No partials in global memory, no ticket, no second launch:
// 12.3 us vs 15.3-17.6 us for the ticket scheme on (4096 4096) (gpurun_out/redbench6.log).
constexpr int kColsClusterSize = 8;
template <class M>
__global__ void __cluster_dims__(1, kColsClusterSize, 1) __launch_bounds__(256)
reduce_cols_cluster_kernel(const __grid_constant__ ColArgs a, const Fin fin) {
  constexpr int LX = 16, LY = 16, U = 4;
  __shared__ float4 sm[LY][LX];
  __shared__ float4 res[LX];
  pdl_prologue();
  cg::cluster_group cluster = cg::this_cluster();
  const int tx = threadIdx.x % LX, ty = threadIdx.x / LX;
  const int64_t b = blockIdx.z;
  const int64_t col = ((int64_t)blockIdx.x * LX + tx) * 4;
  const bool live = col < a.K;
  const int64_t r0 = (int64_t)blockIdx.y * a.rows_per_split;
  int64_t r1 = r0 + a.rows_per_split;
  if (r1 > a.R) r1 = a.R;
  const float* base = a.src + b * a.sB_src + (live ? col : 0);
  float* outp = a.dst + b * a.sB_dst + col;
  const bool writer = live && ty == 0 && blockIdx.y == 0;
  float4 d_old = make_float4(0.f, 0.f, 0.f, 0.f);
  if (writer && fin.reads_dst()) d_old = *reinterpret_cast<const float4*>(outp);
  float4 acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = make_float4(M::id(), M::id(), M::id(), M::id());
  int64_t r = r0 + ty;
  for (; r + LY * (U - 1) < r1; r += LY * U) {
    float4 x[U];
#pragma unroll
    for (int k = 0; k < U; ++k) x[k] = *reinterpret_cast<const float4*>(base + (r + LY * k) * a.sR_src);
#pragma unroll
    for (int k = 0; k < U; ++k) {
      acc[k].x = M::op(acc[k].x, x[k].x); acc[k].y = M::op(acc[k].y, x[k].y);
      acc[k].z = M::op(acc[k].z, x[k].z); acc[k].w = M::op(acc[k].w, x[k].w);
    }
  }
  for (; r < r1; r += LY) {
    const float4 x = *reinterpret_cast<const float4*>(base + r * a.sR_src);
    acc[0].x = M::op(acc[0].x, x.x); acc[0].y = M::op(acc[0].y, x.y); acc[0].z = M::op(acc[0].z, x.z); acc[0].w = M::op(acc[0].w, x.w);
  }
  float4 t = acc[0];
#pragma unroll
  for (int k = 1; k < U; ++k) { t.x = M::op(t.x, acc[k].x); t.y = M::op(t.y, acc[k].y); t.z = M::op(t.z, acc[k].z); t.w = M::op(t.w, acc[k].w); }
  sm[ty][tx] = t;
  __syncthreads();
  if (ty == 0) {
#pragma unroll
    for (int k = 1; k < LY; ++k) { const float4 o = sm[k][tx]; t.x = M::op(t.x, o.x); t.y = M::op(t.y, o.y); t.z = M::op(t.z, o.z); t.w = M::op(t.w, o.w); }
    res[tx] = t;
  }
  cluster.sync();
  if (blockIdx.y == 0 && ty == 0) {   // cluster rank 0 (cluster spans blockIdx.y 0..7: grid.y == 8)
    float4 o[kColsClusterSize];
#pragma unroll
    for (int k = 1; k < kColsClusterSize; ++k) o[k] = *cluster.map_shared_rank(&res[tx], k);
#pragma unroll
    for (int k = 1; k < kColsClusterSize; ++k) { t.x = M::op(t.x, o[k].x); t.y = M::op(t.y, o[k].y); t.z = M::op(t.z, o[k].z); t.w = M::op(t.w, o[k].w); }
    if (writer) {
      d_old.x = fin(d_old.x, t.x); d_old.y = fin(d_old.y, t.y); d_old.z = fin(d_old.z, t.z); d_old.w = fin(d_old.w, t.w);
      *reinterpret_cast<float4*>(outp) = d_old;
    }
  }
  cluster.sync();   // peers keep their strips alive until rank 0 has read them
}

template <class M>
__global__ void __launch_bounds__(256) reduce_cols_final_kernel(const __grid_constant__ ColArgs a, const Fin fin) {
  const int64_t c = (int64_t)blockIdx.x * 256 + threadIdx.x;
  const int64_t b = blockIdx.y;
  if (c >= a.K) return;
  const float* p = a.partial + b * a.splits * a.K + c;
  float acc = M::id();
  for (uint32_t s = 0; s < a.splits; ++s) acc = M::op(acc, p[(int64_t)s * a.K]);
  float* d = a.dst + b * a.sB_dst + c;
  fin.store(d, acc);
}

// ------------------------------------------------------------------ generic
struct GenRedArgs {
  const float* src;
  float* dst;
  int nkept, nred;
  int64_t kshape[kMaxRank], kstride_src[kMaxRank], kstride_dst[kMaxRank];
  int64_t rshape[kMaxRank], rstride_src[kMaxRank];
  int64_t nout, nredel;
};

template <class M>
__global__ void __launch_bounds__(256) reduce_gen_kernel(const __grid_constant__ GenRedArgs a, const Fin fin) {
  const int lane = threadIdx.x & 31;
  const int64_t warps_total = (int64_t)gridDim.x * 8;
  for (int64_t o = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); o < a.nout; o += warps_total) {
    int64_t rem = o, so = 0, dofs = 0;
    for (int d = a.nkept - 1; d >= 0; --d) {
      const int64_t q = rem / a.kshape[d], r = rem - q * a.kshape[d];
      rem = q; so += r * a.kstride_src[d]; dofs += r * a.kstride_dst[d];
    }
    float acc = M::id();
    for (int64_t i = lane; i < a.nredel; i += 32) {
      int64_t rr = i, off = so;
      for (int d = a.nred - 1; d >= 0; --d) {
        const int64_t q = rr / a.rshape[d], r = rr - q * a.rshape[d];
        rr = q; off += r * a.rstride_src[d];
      }
      acc = M::op(acc, a.src[off]);
    }
    acc = warp_reduce<M>(acc);
    if (lane == 0) fin.store(a.dst + dofs, acc);
  }
}

// ------------------------------------------------------------------ host dispatch
static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

bool has_reduction_axes(const wf2cu_view* dst) {
  for (int i = 0; i < dst->rank; ++i)
    if (dst->stride[i] == 0 && dst->shape[i] > 1) return true;
  return false;
}

// Launch shape of the flat kernel (defaults measured on B200, scripts/flat_variants.py -> profiles/r02j_flat_variants.log;
// WF2CU_FLAT_* for A/B runs).  (4096 4096): 4 stages 14.2 us, 3 stages 12.9 us (a 48 KiB ring lets FOUR CTAs share an SM, so
// under programmatic dependent launch the whole next grid is resident and past its barrier set-up when this one drains),
// 2 stages 13.4; 3 or 4 CTAs per SM, an ordinary launch, or letting the dependents in late are all slower.
//   stages x 16 KiB of ring per CTA, CTAs per SM, pdl: 0 = ordinary launch, 1 = programmatic dependent launch with the
//   dependents let in at once, 2 = ... let in when this CTA has finished streaming
// The carve-out matters for the kernels AROUND this one: left to the driver, a 48 KiB ring asks for the 196 KB configuration (four
// CTAs would fit), its neighbours in a compiled sequence use almost no shared memory, and the SMs reconfigure at both boundaries —
// the 12-kernel sweep step took 0.2875 ms instead of 0.2808 although this kernel alone got 1 us faster.  44 % (the 100 KB
// configuration: the two resident CTAs' rings) keeps both: step 0.2806 ms, this kernel 13.2 us (profiles/r02z_flat_carveout.log).
static int flat_carveout() {
  const char* e = getenv("WF2CU_FLAT_CARVEOUT");
  return e ? atoi(e) : 44;
}
struct FlatTune { int stages, ctas_per_sm, pdl; };
static const FlatTune& flat_tune() {
  static FlatTune t = {0, 0, 0};
  if (!t.stages) {
    auto env = [](const char* k, int d, int lo, int hi) { const char* e = getenv(k); int v = e ? atoi(e) : d; return v < lo ? lo : (v > hi ? hi : v); };
    t.ctas_per_sm = env("WF2CU_FLAT_CTAS", 2, 1, 4);
    t.pdl = env("WF2CU_FLAT_PDL", 1, 0, 2);
    t.stages = env("WF2CU_FLAT_STAGES", 3, 2, kFlatStages);
  }
  return t;
}

template <class M>
static int reduce_impl(const wf2cu_view* dst, const wf2cu_view* src, Fin fin, const char* who) {
  const wf2cu_view* views[2] = {dst, src};
  Plan p;
  int rc = build_plan(views, 2, &p, who);
  if (rc) return rc;
  if (p.numel == 0) return 0;
  State& s = S();
  float* dptr = fptr(dst);
  const float* sptr = fptr(src);
  // classify axes
  bool red[kMaxRank];
  int nred = 0;
  for (int d = 0; d < p.rank; ++d) { red[d] = (p.stride[0][d] == 0 && p.shape[d] > 1); nred += red[d]; }
  const int last = p.rank - 1;

  // ---- rows: exactly one reduced axis, innermost, src contiguous there; <= 4 kept axes
  if (nred == 1 && red[last] && p.stride[1][last] == 1 && p.rank - 1 <= 4) {
    RowArgs a;
    memset(&a, 0, sizeof(a));
    a.src = sptr; a.dst = dptr;
    a.nkept = p.rank - 1;
    int64_t rows = 1;
    bool vec = al16(sptr);
    for (int d = 0; d < last; ++d) {
      a.kshape[d] = (uint32_t)p.shape[d];
      a.kstride_src[d] = p.stride[1][d];
      a.kstride_dst[d] = p.stride[0][d];
      rows *= p.shape[d];
      if (p.stride[1][d] % 4 != 0) vec = false;
    }
    a.cols = p.shape[last];
    a.vec = vec ? 1 : 0;
    if (rows == 1 && vec && a.cols >= (int64_t(1) << 18)) {
      // flat: bulk-async ring, 2 persistent CTAs per SM
      static bool attr_set[2] = {false, false};
      auto kfn = reduce_flat_bulk_kernel<M>;
      constexpr int which = std::is_same<M, MSum>::value ? 0 : 1;
      if (!attr_set[which]) {
        WF2_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, kFlatStages * kFlatChunk));
        WF2_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributePreferredSharedMemoryCarveout, flat_carveout()));
        attr_set[which] = true;
      }
      const FlatTune& ft = flat_tune();
      const int64_t nchunks = (a.cols * 4) / kFlatChunk;
      int64_t grid = (int64_t)s.sm_count * ft.ctas_per_sm;
      if (grid > nchunks) grid = nchunks;
      rc = ensure_scratch((size_t)grid * sizeof(float));
      if (rc) return rc;
      FlatArgs f{sptr, dptr, (float*)scratch_base(), counters_base(), a.cols, ft.stages, ft.pdl == 2 ? 1 : 0};
      if (ft.pdl) WF2_CUDA(launch_pdl(kfn, dim3((unsigned)grid), dim3(kFlatThreads), ft.stages * kFlatChunk, s.stream, f, fin));
      else { kfn<<<(unsigned)grid, kFlatThreads, ft.stages * kFlatChunk, s.stream>>>(f, fin); }
      return post_launch(who);
    }
    if (rows < (int64_t(1) << 31)) {
      // register-tile kernel: all loads of a row tile issued up front
      if (a.vec && a.cols >= 4 && a.cols <= 8192 && rows * a.cols >= (int64_t(1) << 16)) {
        const int64_t v1 = a.cols >> 2;
        a.chunks = 1; a.chunk_len = a.cols;
        if (v1 <= 256) {
          WF2_CUDA(launch_pdl(reduce_rows_tile_kernel<M, 8>, dim3((unsigned)((rows + 7) / 8)), dim3(256), 0, s.stream, a, fin, (uint32_t)rows));
        } else if (v1 <= 512) {
          WF2_CUDA(launch_pdl(reduce_rows_tile_kernel<M, 4>, dim3((unsigned)((rows + 3) / 4)), dim3(256), 0, s.stream, a, fin, (uint32_t)rows));
        } else if (v1 <= 1024) {
          WF2_CUDA(launch_pdl(reduce_rows_tile_kernel<M, 2>, dim3((unsigned)((rows + 1) / 2)), dim3(256), 0, s.stream, a, fin, (uint32_t)rows));
        } else {
          WF2_CUDA(launch_pdl(reduce_rows_tile_kernel<M, 1>, dim3((unsigned)rows), dim3(256), 0, s.stream, a, fin, (uint32_t)rows));
        }
        return post_launch(who);
      }
      const int64_t fill = (int64_t)s.sm_count * 16;   // warps needed to keep every SM busy
      if ((a.cols <= 1024 && rows >= 64) || (rows >= fill && a.cols <= 65536)) {
        a.chunks = 1; a.chunk_len = a.cols;
        reduce_rows_warp_kernel<M><<<(unsigned)((rows + 7) / 8), 256, 0, s.stream>>>(a, fin, (uint32_t)rows);
        return post_launch(who);
      }
      // split long rows when there are too few of them to fill 148 SMs; a chunk is a whole number
      // of 8192-float units (256 threads x 8 float4)
      uint32_t chunks = 1;
      int64_t chunk_len = a.cols;
      const int64_t target_ctas = (int64_t)s.sm_count * 8;
      if (rows < target_ctas && a.cols >= 8192 && rows <= 4096) {
        const int64_t units = (a.cols + 8191) / 8192;
        int64_t want = (target_ctas + rows - 1) / rows;
        if (want > units) want = units;
        const int64_t upc = (units + want - 1) / want;   // units per chunk
        chunk_len = upc * 8192;
        chunks = (uint32_t)((a.cols + chunk_len - 1) / chunk_len);
      }
      a.chunks = chunks; a.chunk_len = chunk_len;
      if (chunks > 1) {
        rc = ensure_scratch((size_t)rows * chunks * sizeof(float));
        if (rc) return rc;
        a.partial = (float*)scratch_base();
        a.counters = counters_base();
      }
      reduce_rows_kernel<M, 256><<<(unsigned)(rows * chunks), 256, 0, s.stream>>>(a, fin);
      return post_launch(who);
    }
  }
  // ---- cols: [B,] R, K with K contiguous in both, R reduced, B kept
  if (nred == 1 && p.rank >= 2 && p.rank <= 3 && red[p.rank - 2] && p.stride[0][last] == 1 && p.stride[1][last] == 1) {
    ColArgs a;
    memset(&a, 0, sizeof(a));
    a.src = sptr; a.dst = dptr;
    a.K = p.shape[last]; a.R = p.shape[last - 1];
    a.sR_src = p.stride[1][last - 1];
    if (p.rank == 3) { a.B = p.shape[0]; a.sB_src = p.stride[1][0]; a.sB_dst = p.stride[0][0]; }
    else { a.B = 1; }
    bool vec = al16(sptr) && al16(dptr) && (a.K % 4 == 0) && (a.sR_src % 4 == 0) && (a.sB_src % 4 == 0) && (a.sB_dst % 4 == 0);
    a.vec = vec ? 1 : 0;
    {
      // cluster path: 64-column strips x 8 CTAs along R, partial strips folded through DSMEM
      const int64_t strips = (a.K + 63) / 64;
      if (vec && a.R >= 512 && a.B < 65536 && strips * kColsClusterSize * a.B >= (int64_t)s.sm_count * 2 && strips < (int64_t(1) << 31)) {
        a.splits = kColsClusterSize;
        a.rows_per_split = (a.R + kColsClusterSize - 1) / kColsClusterSize;
        dim3 grid((unsigned)strips, kColsClusterSize, (unsigned)a.B);
        WF2_CUDA(launch_pdl(reduce_cols_cluster_kernel<M>, grid, dim3(256), 0, s.stream, a, fin));
        return post_launch(who);
      }
    }
    if (a.B == 1 && a.K <= 64 && a.sR_src == a.K && a.R >= 256) {
      // narrow and tall: one flat contiguous run
      const int64_t L = 256 / a.K;
      int64_t splits = (a.R + 4 * L * 2 - 1) / (4 * L * 2);          // two unrolled steps (8 loads) per thread: latency-bound otherwise
      if (splits > (int64_t)s.sm_count * 2) splits = (int64_t)s.sm_count * 2;
      int64_t rps = (a.R + splits - 1) / splits;
      splits = (a.R + rps - 1) / rps;
      a.splits = (uint32_t)splits; a.rows_per_split = rps;
      if (splits > 1) {
        rc = ensure_scratch((size_t)splits * a.K * sizeof(float));
        if (rc) return rc;
        a.partial = (float*)scratch_base();
        a.counters = counters_base();
      }
      reduce_cols_narrow_kernel<M><<<(unsigned)splits, 256, 0, s.stream>>>(a, fin);
      return post_launch(who);
    }
    const int64_t colblocks = (a.K + 127) / 128;
    const int64_t target_ctas = (int64_t)s.sm_count * 8;
    int64_t splits = 1;
    if (colblocks * a.B < target_ctas) {
      splits = (target_ctas + colblocks * a.B - 1) / (colblocks * a.B);
      const int64_t maxs = (a.R + 31) / 32;  // >= 32 rows per split
      if (splits > maxs) splits = maxs;
      if (splits < 1) splits = 1;
    }
    int64_t rps = (a.R + splits - 1) / splits;
    splits = (a.R + rps - 1) / rps;
    a.splits = (uint32_t)splits; a.rows_per_split = rps;
    if (a.B < 65536 && splits < 65536) {
      const bool fused2 = splits > 1 && colblocks * a.B <= 4096;
      if (splits > 1) {
        rc = ensure_scratch((size_t)a.B * splits * a.K * sizeof(float));
        if (rc) return rc;
        a.partial = (float*)scratch_base();
        a.counters = fused2 ? counters_base() : nullptr;
      }
      dim3 grid((unsigned)colblocks, (unsigned)splits, (unsigned)a.B);
      reduce_cols_kernel<M><<<grid, 256, 0, s.stream>>>(a, fin);
      rc = post_launch(who);
      if (rc) return rc;
      if (splits > 1 && !fused2) {
        dim3 g2((unsigned)((a.K + 255) / 256), (unsigned)a.B);
        reduce_cols_final_kernel<M><<<g2, 256, 0, s.stream>>>(a, fin);
        rc = post_launch(who);
      }
      return rc;
    }
  }
  // ---- generic
  GenRedArgs g;
  memset(&g, 0, sizeof(g));
  g.src = sptr; g.dst = dptr;
  g.nout = 1; g.nredel = 1;
  for (int d = 0; d < p.rank; ++d) {
    if (red[d]) {
      g.rshape[g.nred] = p.shape[d]; g.rstride_src[g.nred] = p.stride[1][d]; g.nredel *= p.shape[d]; ++g.nred;
    } else {
      g.kshape[g.nkept] = p.shape[d]; g.kstride_src[g.nkept] = p.stride[1][d]; g.kstride_dst[g.nkept] = p.stride[0][d];
      g.nout *= p.shape[d]; ++g.nkept;
    }
  }
  int64_t grid = (g.nout + 7) / 8;
  const int64_t cap = (int64_t)s.sm_count * 16;
  if (grid > cap) grid = cap;
  reduce_gen_kernel<M><<<(unsigned)grid, 256, 0, s.stream>>>(g, fin);
  return post_launch(who);
}

int reduce_binary(int op, const wf2cu_view* dst, const wf2cu_view* src) {
  switch (op) {
    case WF2CU_ADD: return reduce_impl<MSum>(dst, src, Fin{FIN_ADD, 0.f}, "wf2cu_ew_binary(add,reduce)");
    case WF2CU_SUB: return reduce_impl<MSum>(dst, src, Fin{FIN_SUB, 0.f}, "wf2cu_ew_binary(sub,reduce)");
    case WF2CU_MUL: return reduce_impl<MProd>(dst, src, Fin{FIN_MUL, 0.f}, "wf2cu_ew_binary(mul,reduce)");
    case WF2CU_DIV: return reduce_impl<MProd>(dst, src, Fin{FIN_DIV, 0.f}, "wf2cu_ew_binary(div,reduce)");
    case WF2CU_COPY: {
      // sequential `x[0] = y[i]` leaves the LAST element along every reduced axis
      wf2cu_view d = *dst, s = *src;
      for (int i = 0; i < d.rank; ++i)
        if (d.stride[i] == 0 && d.shape[i] > 1) {
          s.offset += s.stride[i] * (s.shape[i] - 1);
          s.shape[i] = 1; d.shape[i] = 1;
        }
      return wf2cu_ew_binary(WF2CU_COPY, &d, &s);
    }
    default: return set_error(WF2CU_ERR_INVALID, "wf2cu_ew_binary: unknown op %d", op);
  }
}

}  // namespace wf2

using namespace wf2;

extern "C" {

int wf2cu_reduce_sum(const wf2cu_view* x, const wf2cu_view* out_bcast, int accumulate, float divisor) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(x && out_bcast, "wf2cu_reduce_sum: null view");
  WF2_CHECK(x->rank == out_bcast->rank, "wf2cu_reduce_sum: rank mismatch (%d vs %d)", x->rank, out_bcast->rank);
  // accept out given with extent-1 axes instead of (:broadcast n): widen them to stride 0
  wf2cu_view o = *out_bcast;
  for (int i = 0; i < o.rank; ++i)
    if (o.shape[i] == 1 && x->shape[i] != 1) { o.shape[i] = x->shape[i]; o.stride[i] = 0; }
  Fin fin{accumulate ? FIN_ADD : FIN_SET, divisor};
  if (!has_reduction_axes(&o)) {
    // nothing to reduce: out <- [out +] x [/ divisor]
    int rc = accumulate ? wf2cu_ew_binary(WF2CU_ADD, &o, x) : wf2cu_ew_binary(WF2CU_COPY, &o, x);
    if (rc) return rc;
    if (divisor != 0.f) return wf2cu_ew_scalar(WF2CU_DIV, &o, divisor);
    return 0;
  }
  return reduce_impl<MSum>(&o, x, fin, "wf2cu_reduce_sum");
}

}  // extern "C"

// ------------------------------------------------------------------ allocation-free `sum :axis t`
// wf2cu_sum_all_noalloc: reduce_flat_bulk_kernel<MSum> fed from a file-scope device array instead of the cudaMalloc'ed
// scratch of runtime.cu, for callers that may not allocate (graph capture, serving loops).  Layout of the array:
// [0, kNoAllocSlots) one fp32 partial per CTA, then one 256-byte line whose first word is the last-block ticket.
// Every GPU that loads the library holds these kNoAllocBytes until exit.  All calls share the array: calls that
// overlap on the GPU (two streams, two host threads) race on the partials and on the ticket.
namespace wf2 {
constexpr int kNoAllocSlots = 1024;                                   // CTAs a launch may use
constexpr size_t kNoAllocBytes = kNoAllocSlots * sizeof(float) + 256;
constexpr int kNoAllocMaxDevices = 64;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// In host code `g_noalloc_scratch` is the host-side stand-in of the symbol, not a device address: the address the
// kernel gets is looked up once per device (with that device current) and kept here.
static unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {};

static int noalloc_setup(int dev, unsigned char** out) {
  if (dev < 0 || dev >= kNoAllocMaxDevices)
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_sum_all_noalloc: device ordinal %d outside 0..%d", dev, kNoAllocMaxDevices - 1);
  if (!g_noalloc_ptr[dev]) {
    unsigned char* p = nullptr;
    WF2_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    WF2_CUDA(cudaFuncSetAttribute(reduce_flat_bulk_kernel<MSum>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFlatStages * kFlatChunk));
    WF2_CUDA(cudaFuncSetAttribute(reduce_flat_bulk_kernel<MSum>, cudaFuncAttributePreferredSharedMemoryCarveout, flat_carveout()));
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace wf2

extern "C" {

int wf2cu_sum_all_noalloc(const void* x, int64_t x_off, int64_t n, void* out, int64_t out_off, int accumulate, float divisor) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(x && out, "wf2cu_sum_all_noalloc: null pointer");
  WF2_CHECK(n >= 1, "wf2cu_sum_all_noalloc: n = %lld (need at least one element)", (long long)n);
  if (n > (int64_t)WF2CU_NOALLOC_MAX_ELEMS)
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_sum_all_noalloc: n = %lld exceeds the maximum of %lld elements this entry point "
                     "accepts (its scratch is a fixed %zu-byte device array); use wf2cu_reduce_sum",
                     (long long)n, (long long)WF2CU_NOALLOC_MAX_ELEMS, kNoAllocBytes);
  const float* sptr = reinterpret_cast<const float*>(x) + x_off;
  float* dptr = reinterpret_cast<float*>(out) + out_off;
  WF2_CHECK(al16(sptr), "wf2cu_sum_all_noalloc: x + x_off must be 16-byte aligned");
  State& s = S();
  int dev = -1;
  WF2_CUDA(cudaGetDevice(&dev));
  unsigned char* scratch = nullptr;
  int rc = noalloc_setup(dev, &scratch);
  if (rc) return rc;
  float* partial = reinterpret_cast<float*>(scratch);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + kNoAllocSlots * sizeof(float));
  // same grid rule as wf2cu_reduce_sum's flat path (2 persistent CTAs per SM), so both give the same bits
  const int64_t nchunks = (n * 4) / kFlatChunk;
  const FlatTune& ft = flat_tune();
  int64_t grid = (int64_t)s.sm_count * ft.ctas_per_sm;
  if (grid > nchunks) grid = nchunks;
  if (grid < 1) grid = 1;               // shorter than one chunk: CTA 0's ragged-tail loop does all of it
  if (grid > kNoAllocSlots) grid = kNoAllocSlots;
  // the ticket must start at zero whatever an earlier (possibly aborted) call left behind; the partials are
  // overwritten before they are read
  WF2_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), s.stream));
  FlatArgs f{sptr, dptr, partial, ticket, n, ft.stages, 0};
  Fin fin{accumulate ? FIN_ADD : FIN_SET, divisor};
  reduce_flat_bulk_kernel<MSum><<<(unsigned)grid, kFlatThreads, ft.stages * kFlatChunk, s.stream>>>(f, fin);
  return post_launch("wf2cu_sum_all_noalloc");
}

}  // extern "C"
