// gemm.cu — MatmulNode entry point and the exact-fp32 SIMT path.
//
// The reference runs `!matmul` as one Fortran-ABI `sgemm_` per batch index, alpha=1 beta=0, with
// the row-major argument swap and the `LazyTransposeNode` flags turned into trans characters
// (`source/backends/cpu/matrix-ops.lisp:12-237`, `blas-functions.lisp:34-38`).  Here the batch loop
// is a grid dimension.  Two device paths:
//   * gemm_tc.cu : tcgen05/TMEM tensor-core kernel fed by TMA, 3xTF32 error-compensated split so
//                  the result stays fp32-faithful (SURVEY §7 hard part c) — large aligned shapes.
//   * this file  : register-tiled fp32 FMA kernel (128x128x16 tile, 8x8 per thread) for everything
//                  the tensor path does not take (tiny / unaligned shapes such as the 10-wide
//                  classifier head).
#include <stdlib.h>
#include "common.cuh"
#include "math.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace wf2 {

// implemented in gemm_tc.cu; returns 1 when it handled the problem, 0 when it declines, <0 error
int gemm_tc_try(int ta, int tb, int64_t m, int64_t n, int64_t k, const float* A, int64_t lda, const float* B,
                int64_t ldb, float* C, int64_t ldc, int64_t batch, int64_t sa, int64_t sb, int64_t sc, int mode, const float* bias, int act,
                const float* res, int64_t ldres);
// gemm_tc.cu: one-launch tcgen05 kernel for small problems (in-kernel hi/lo split, cluster split-K, fused epilogue)
int gemm_raw_try(int ta, int tb, int64_t m, int64_t n, int64_t k, const float* A, int64_t lda, const float* B, int64_t ldb,
                 float* C, int64_t ldc, int64_t batch, int64_t sa, int64_t sb, int64_t sc, const float* bias, int act, const float* res,
                 int64_t ldres);
bool gemm_raw_preferred(int tb, int64_t m, int64_t n, int64_t k, int64_t batch, double a_share, double b_share);
double gemm_tc_prepass_share(const float* p, int64_t ld, int64_t bstride, int64_t nbatch, bool kmajor);
int gemm_tc_const_declare(const void* ptr, size_t bytes, int on);
void gemm_tc_const_stats(int64_t* entries, int64_t* bytes);
void gemm_tc_set_correction(int kind);
int gemm_tc_correction();
int gemm_tc_fused_try(int ta, int tb, int64_t m, int64_t n, int64_t k, const float* A, int64_t lda, const float* B, int64_t ldb,
                      float* C, int64_t ldc, int64_t batch, int64_t sa, int64_t sb, int64_t sc, const float* bias, int act,
                      const float* mask, int64_t ldmask);

constexpr int BM = 128, BN = 128, BK = 16;

struct GemmArgs {
  const float* A; const float* B; float* C;
  int64_t m, n, k, lda, ldb, ldc, sa, sb, sc;
  int ta, tb;
  // fused epilogue (wf2cu_gemm_ex): v = acc [+ bias[col]] ; [v = v * (v > 0)] ; [v = v * (mask[row, col] > 0)]
  const float* bias; const float* mask; int64_t ldmask; int act;
  const float* res = nullptr; int64_t ldres = 0;     // [v = v + res[row, col]]  (a residual connection riding on the projection)
};

// `!relu` is x * (where x > 0 then 1 else 0) (nn/activation.lisp:11-26): a true multiply, so a negative
// x gives -0.0 exactly as the reference's MulNode does.
__device__ __forceinline__ float epilogue(const GemmArgs& g, float v, int64_t gm, int64_t gn) {
  if (g.bias) v = __fadd_rn(v, g.bias[gn]);
  if (g.act == 1) v = __fmul_rn(v, v > 0.f ? 1.f : 0.f);
  else if (g.act == 2) v = gelu_tanh_ref(v);
  if (g.res) v = __fadd_rn(g.res[gm * g.ldres + gn], v);
  if (g.mask) v = __fmul_rn(v, g.mask[gm * g.ldmask + gn] > 0.f ? 1.f : 0.f);
  return v;
}

// the same epilogue as its own pass, for GEMMs that ran on the tensor-core path
__global__ void __launch_bounds__(256) gemm_epilogue_kernel(const __grid_constant__ GemmArgs g) {
  const int64_t total = g.m * g.n;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t gm = i / g.n, gn = i - gm * g.n;
    float* c = g.C + gm * g.ldc + gn;
    *c = epilogue(g, *c, gm, gn);
  }
}

__global__ void __launch_bounds__(256) sgemm_simt_kernel(const __grid_constant__ GemmArgs g) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int64_t b = blockIdx.z;
  const float* A = g.A + b * g.sa;
  const float* B = g.B + b * g.sb;
  float* C = g.C + b * g.sc;
  const int64_t m0 = (int64_t)blockIdx.y * BM, n0 = (int64_t)blockIdx.x * BN;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, each 8x8 outputs (2x2 blocks of 4x4)
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  for (int64_t k0 = 0; k0 < g.k; k0 += BK) {
    // ---- stage A tile (BM x BK) as As[k][m]
    if (!g.ta) {
      // A[m, k] row-major: k contiguous -> threads run along k
#pragma unroll
      for (int it = 0; it < (BM * BK) / 256; ++it) {
        const int e = tid + it * 256;
        const int kk = e & (BK - 1), mm = e >> 4;
        const int64_t gm = m0 + mm, gk = k0 + kk;
        As[kk][mm] = (gm < g.m && gk < g.k) ? A[gm * g.lda + gk] : 0.f;
      }
    } else {
      // stored transposed: A^T is k x m, m contiguous
#pragma unroll
      for (int it = 0; it < (BM * BK) / 256; ++it) {
        const int e = tid + it * 256;
        const int mm = e & (BM - 1), kk = e >> 7;
        const int64_t gm = m0 + mm, gk = k0 + kk;
        As[kk][mm] = (gm < g.m && gk < g.k) ? A[gk * g.lda + gm] : 0.f;
      }
    }
    // ---- stage B tile (BK x BN) as Bs[k][n]
    if (!g.tb) {
#pragma unroll
      for (int it = 0; it < (BN * BK) / 256; ++it) {
        const int e = tid + it * 256;
        const int nn = e & (BN - 1), kk = e >> 7;
        const int64_t gn = n0 + nn, gk = k0 + kk;
        Bs[kk][nn] = (gn < g.n && gk < g.k) ? B[gk * g.ldb + gn] : 0.f;
      }
    } else {
      // stored transposed: B^T is n x k, k contiguous
#pragma unroll
      for (int it = 0; it < (BN * BK) / 256; ++it) {
        const int e = tid + it * 256;
        const int kk = e & (BK - 1), nn = e >> 4;
        const int64_t gn = n0 + nn, gk = k0 + kk;
        Bs[kk][nn] = (gn < g.n && gk < g.k) ? B[gn * g.ldb + gk] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[8], bb[8];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[kk][64 + tx * 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      bb[0] = b0.x; bb[1] = b0.y; bb[2] = b0.z; bb[3] = b0.w; bb[4] = b1.x; bb[5] = b1.y; bb[6] = b1.z; bb[7] = b1.w;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t gm = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (gm >= g.m) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int64_t gn = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (gn < g.n) C[gm * g.ldc + gn] = epilogue(g, acc[i][j], gm, gn);
    }
  }
}

// ------------------------------------------------------------------ small / skinny problems: 64x64 tiles, cluster split-K
// The MLP training step (BASELINE config 4) is made of GEMMs with 4..64 output tiles — (1024 x 256 x 784),
// (256 x 256 x 1024), (10 x 256 x 1024) ... — which left 140 of 148 SMs idle in the kernels above (170 us
// for the 10 x 256 x 1024 weight gradient on 2 CTAs, profiles/r01_mlp_launches_before.md).  Here a tile
// is 64 x 64 (4 x 4 per thread) and the K axis is split over a THREAD-BLOCK CLUSTER of up to 8 CTAs:
// every CTA accumulates its K slice, parks the partial tile in shared memory, and after one
// cluster barrier each CTA sums one strip of the tile over all peers through distributed shared
// memory, in slice order (deterministic), and writes C.  No partials in global memory, no atomics.
constexpr int SBM = 64, SBN = 64, SBK = 16, SSTAGES = 4;
constexpr int SLD = SBM + 4;                       // padded row of a staged k-slice
constexpr int STAGE_FLOATS = 2 * SBK * SLD;        // A slice + B slice

__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc, bool valid) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  const int n = valid ? 4 : 0;                     // src-size 0: the 4 bytes are zero-filled, nothing is read
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
}

__global__ void __launch_bounds__(256) sgemm_splitk_kernel(const __grid_constant__ GemmArgs g, const int ksplit, const int64_t kchunk) {
  // 4-stage cp.async ring of k-slices (A as [k][m], B as [k][n]); after the main loop the same storage
  // holds this CTA's 64 x 64 partial tile for the cluster fold.
  __shared__ __align__(16) float ring[SSTAGES * STAGE_FLOATS];
  static_assert(SSTAGES * STAGE_FLOATS >= SBM * SBN, "partial tile must fit the ring");
  const int ks = (int)(blockIdx.z % ksplit);
  const int64_t b = blockIdx.z / ksplit;
  const float* A = g.A + b * g.sa;
  const float* B = g.B + b * g.sb;
  float* C = g.C + b * g.sc;
  const int64_t m0 = (int64_t)blockIdx.y * SBM, n0 = (int64_t)blockIdx.x * SBN;
  const int64_t kbeg = (int64_t)ks * kchunk;
  const int64_t kend = kbeg + kchunk < g.k ? kbeg + kchunk : g.k;
  const int nkb = kend > kbeg ? (int)((kend - kbeg + SBK - 1) / SBK) : 0;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;           // 16 x 16 threads, 4 x 4 outputs each
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  // Each thread copies 4 elements of A and 4 of B per k-slice, 4 bytes at a time (no alignment
  // demands on lda/ldb/offsets), coalesced along the operand's contiguous axis.
  const float* pa[4]; const float* pb[4];
  int oa[4], ob[4], ka[4], kb_[4];
  bool va[4], vb[4];
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    const int e = tid + it * 256;
    int kk, mm;
    if (!g.ta) { kk = e & (SBK - 1); mm = e >> 4; } else { mm = e & (SBM - 1); kk = e >> 6; }
    const int64_t gm = m0 + mm;
    va[it] = gm < g.m; ka[it] = kk; oa[it] = kk * SLD + mm;
    pa[it] = va[it] ? (g.ta ? A + (kbeg + kk) * g.lda + gm : A + gm * g.lda + kbeg + kk) : A;
    int kb, nn;
    if (!g.tb) { nn = e & (SBN - 1); kb = e >> 6; } else { kb = e & (SBK - 1); nn = e >> 4; }
    const int64_t gn = n0 + nn;
    vb[it] = gn < g.n; kb_[it] = kb; ob[it] = SBK * SLD + kb * SLD + nn;
    pb[it] = vb[it] ? (g.tb ? B + gn * g.ldb + kbeg + kb : B + (kbeg + kb) * g.ldb + gn) : B;
  }
  const int64_t stepa = g.ta ? (int64_t)SBK * g.lda : SBK, stepb = g.tb ? SBK : (int64_t)SBK * g.ldb;
  auto issue = [&](int blk) {
    float* st = ring + (blk % SSTAGES) * STAGE_FLOATS;
    const int64_t k0 = kbeg + (int64_t)blk * SBK;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const bool oka = va[it] && k0 + ka[it] < kend, okb = vb[it] && k0 + kb_[it] < kend;
      cp_async4(st + oa[it], oka ? pa[it] + blk * stepa : A, oka);
      cp_async4(st + ob[it], okb ? pb[it] + blk * stepb : B, okb);
    }
  };
#pragma unroll
  for (int sgs = 0; sgs < SSTAGES - 1; ++sgs) {
    if (sgs < nkb) issue(sgs);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (int blk = 0; blk < nkb; ++blk) {
    asm volatile("cp.async.wait_group %0;" ::"n"(SSTAGES - 2) : "memory");
    __syncthreads();                                // slice blk has landed; slice blk-1's stage is free
    if (blk + SSTAGES - 1 < nkb) issue(blk + SSTAGES - 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    const float* as = ring + (blk % SSTAGES) * STAGE_FLOATS;
    const float* bs = as + SBK * SLD;
#pragma unroll
    for (int kk = 0; kk < SBK; ++kk) {
      const float4 a4 = *reinterpret_cast<const float4*>(as + kk * SLD + ty * 4);
      const float4 b4 = *reinterpret_cast<const float4*>(bs + kk * SLD + tx * 4);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w}, bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
  }

  if (ksplit == 1) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t gm = m0 + ty * 4 + i;
      if (gm >= g.m) continue;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t gn = n0 + tx * 4 + j;
        if (gn < g.n) C[gm * g.ldc + gn] = epilogue(g, acc[i][j], gm, gn);
      }
    }
    return;
  }
  // ---- split-K: partial tile -> shared memory, fold across the cluster through DSMEM
  cg::cluster_group cluster = cg::this_cluster();
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();                                  // everyone is done reading the ring
  float (*Ps)[SBN] = reinterpret_cast<float (*)[SBN]>(ring);
#pragma unroll
  for (int i = 0; i < 4; ++i)
    *reinterpret_cast<float4*>(&Ps[ty * 4 + i][tx * 4]) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  cluster.sync();
  const int rows_per = SBM / ksplit;                // this CTA folds rows [ks*rows_per, (ks+1)*rows_per)
  for (int e = tid; e < rows_per * (SBN / 4); e += 256) {
    const int r = ks * rows_per + e / (SBN / 4), c4 = (e % (SBN / 4)) * 4;
    float4 t = *reinterpret_cast<const float4*>(cluster.map_shared_rank(&Ps[r][c4], 0));
    for (int p = 1; p < ksplit; ++p) {
      const float4 o = *reinterpret_cast<const float4*>(cluster.map_shared_rank(&Ps[r][c4], p));
      t.x += o.x; t.y += o.y; t.z += o.z; t.w += o.w;
    }
    const int64_t gm = m0 + r, gn = n0 + c4;
    if (gm < g.m) {
      float* cp = C + gm * g.ldc + gn;
      const float v[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (gn + j < g.n) cp[j] = epilogue(g, v[j], gm, gn + j);
    }
  }
  cluster.sync();                                   // nobody leaves while a peer may still read its tile
}

// 0 = not taken.  Picks the cluster split-K kernel when the 128x128 grid would leave most SMs idle.
static int sgemm_small_try(const GemmArgs& g, int64_t batch) {
  const int64_t big_tiles = ((g.m + BM - 1) / BM) * ((g.n + BN - 1) / BN) * batch;
  if (big_tiles >= S().sm_count) return 0;
  const int64_t tiles = ((g.m + SBM - 1) / SBM) * ((g.n + SBN - 1) / SBN);
  if (batch * 8 >= 65536) return 0;
  int ksplit = 1;
  while (ksplit < 8 && tiles * batch * ksplit * 2 <= (int64_t)S().sm_count * 2 && g.k / (ksplit * 2) >= 32) ksplit *= 2;
  int64_t kchunk = (g.k + ksplit - 1) / ksplit;
  kchunk = (kchunk + SBK - 1) / SBK * SBK;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)((g.n + SBN - 1) / SBN), (unsigned)((g.m + SBM - 1) / SBM), (unsigned)(batch * ksplit));
  cfg.blockDim = dim3(256, 1, 1);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = S().stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = (unsigned)ksplit;
  cfg.attrs = attr; cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, sgemm_splitk_kernel, g, ksplit, kchunk);
  if (e != cudaSuccess) return -cuda_fail(e, "cudaLaunchKernelEx(sgemm_splitk_kernel)", __FILE__, __LINE__);
  const int rc = post_launch("wf2cu_gemm(simt split-k)");
  return rc ? -rc : 1;
}

// ------------------------------------------------------------------ skinny shapes (Conv2D on few channels)
// Conv2D lowers to (N*Ho*Wo, C*kh*kw) x (C*kh*kw, C_out): on MNIST-CNN that is 67 600 x 9 x 4 and 8 100 x 100 x 16.
// A 64x64 output tile is then almost empty along n, and the weight-gradient GEMM (4 x 9 outputs over k = 67 600)
// would run on one tile.  Two exact-fp32 kernels cover the two shapes.  Both are latency-bound, not
// bandwidth-bound (a few MB, L2-resident): what matters is enough threads and independent loads in flight —
// a first version with one thread per row (8 K threads on the GPU) ran at IPC 0.28 and took 83 us on 8 100 x 100 x 16
// (profiles/r01c_ncu_cnn_skinny.md).

// (a) C[M, n] = A[M, k] . B^T, B stored (n, k), n <= 32, n*k <= 4096: one thread per OUTPUT element, 256 / NP rows
// per CTA (NP = n rounded up to a power of two).  B sits in shared memory with an odd row stride (the NP threads of
// a row read NP different banks), A is staged in k-slices through shared memory with coalesced global reads and is
// read back as a broadcast.
constexpr int SKN_TK = 64;
template <int NP>
__global__ void __launch_bounds__(256) sgemm_skinny_nt_kernel(const __grid_constant__ GemmArgs g) {
  constexpr int R = 256 / NP;
  __shared__ float Bs[4096 + 64];
  __shared__ float As[R * (SKN_TK + 1)];
  const int t = threadIdx.x;
  const int n = (int)g.n, k = (int)g.k;
  const int kp = k | 1;
  for (int i = t; i < n * k; i += 256) { const int jj = i / k, cc = i - jj * k; Bs[jj * kp + cc] = __ldg(g.B + (int64_t)jj * g.ldb + cc); }
  const int64_t row0 = (int64_t)blockIdx.x * R;
  const int rows = (int)((g.m - row0) < R ? (g.m - row0) : R);
  const int r = t / NP, j = t % NP;
  const bool mine = r < rows && j < n;
  float acc0 = 0.f, acc1 = 0.f;
  for (int k0 = 0; k0 < k; k0 += SKN_TK) {
    const int tk = (k - k0) < SKN_TK ? (k - k0) : SKN_TK;
    __syncthreads();
    const int total = rows * tk;
    for (int i0 = t; i0 < total; i0 += 256 * 4) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        // unconditional (clamped) loads: a guarded load is not hoisted above the previous store and the
        // four round trips would run one after the other
        const int i = min(i0 + u * 256, total - 1);
        const int rr = i / tk, cc = i - rr * tk;
        v[u] = __ldg(g.A + (row0 + rr) * g.lda + k0 + cc);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * 256;
        const int rr = i / tk, cc = i - rr * tk;
        if (i < total) As[rr * (SKN_TK + 1) + cc] = v[u];
      }
    }
    __syncthreads();
    if (mine) {
      const float* ap = As + r * (SKN_TK + 1);
      const float* bp = Bs + j * kp + k0;
      int c = 0;
      for (; c + 1 < tk; c += 2) { acc0 = fmaf(ap[c], bp[c], acc0); acc1 = fmaf(ap[c + 1], bp[c + 1], acc1); }
      if (c < tk) acc0 = fmaf(ap[c], bp[c], acc0);
    }
  }
  if (mine) g.C[(row0 + r) * g.ldc + j] = epilogue(g, acc0 + acc1, row0 + r, j);
}

static int sgemm_skinny_nt_try(const GemmArgs& g, int64_t batch) {
  if (batch != 1 || g.ta || !g.tb) return 0;
  if (g.n > 32 || g.n * g.k > 4096 || g.m < 2048) return 0;
  cudaStream_t st = S().stream;
  if (g.n <= 4) sgemm_skinny_nt_kernel<4><<<(unsigned)((g.m + 63) / 64), 256, 0, st>>>(g);
  else if (g.n <= 8) sgemm_skinny_nt_kernel<8><<<(unsigned)((g.m + 31) / 32), 256, 0, st>>>(g);
  else if (g.n <= 16) sgemm_skinny_nt_kernel<16><<<(unsigned)((g.m + 15) / 16), 256, 0, st>>>(g);
  else sgemm_skinny_nt_kernel<32><<<(unsigned)((g.m + 7) / 8), 256, 0, st>>>(g);
  const int rc = post_launch("wf2cu_gemm(skinny nt)");
  return rc ? -rc : 1;
}

// (b) C[m, n] = A^T . B with A stored (K, m), B stored (K, n), m*n <= 2048, K long (a weight gradient): a reduction
// over K with an m x n accumulator.  CTA c owns a slice of K and walks it in chunks of 64 rows staged through shared
// memory (coalesced reads, 4 per thread in flight).  m*n > 256: thread t owns outputs t, t+256, ... (Q of them);
// m*n <= 256: 256 / (m*n) thread groups share the chunk's rows and meet in shared memory.  CTA partials meet through
// a last-block ticket; every load of the fold is independent of the others.  All orders are fixed: deterministic.
struct SkinnyTnArgs {
  GemmArgs g;
  float* partial; unsigned int* counter;
  int64_t rows_per_cta; int ctas;
  int ks;                 // rows per staged chunk: as many as fit 32 KiB (64 .. 512)
  int dense4;             // lda == m, ldb == n, 16-byte aligned, every chunk a whole number of float4
};

template <int Q>
__global__ void __launch_bounds__(256) sgemm_skinny_tn_kernel(const __grid_constant__ SkinnyTnArgs a) {
  extern __shared__ float smem[];         // As[KS][m] | Bs[KS][n] | red[256]
  __shared__ bool is_last;
  const GemmArgs& g = a.g;
  const int m = (int)g.m, n = (int)g.n, mn = m * n;
  const int SKT_KS = a.ks;
  float* As = smem;
  float* Bs = smem + SKT_KS * m;
  float* red = Bs + SKT_KS * n;
  const int t = threadIdx.x;
  const int G = (Q == 1 && mn <= 256) ? 256 / mn : 1;       // thread groups sharing the rows
  const int grp = (G > 1) ? t / mn : 0;
  const int o0 = (G > 1) ? t % mn : t;
  int ai[Q], bj[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) { const int o = o0 + 256 * q; ai[q] = o < mn ? o / n : 0; bj[q] = o < mn ? o % n : 0; }
  const bool active = (G > 1) ? grp < G : true;
  float acc[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) acc[q] = 0.f;
  const int64_t k0 = (int64_t)blockIdx.x * a.rows_per_cta;
  int64_t k1 = k0 + a.rows_per_cta;
  if (k1 > g.k) k1 = g.k;
  for (int64_t kc = k0; kc < k1; kc += SKT_KS) {
    const int rows = (int)((k1 - kc) < SKT_KS ? (k1 - kc) : SKT_KS);
    __syncthreads();
    {
      // A rows then B rows of the chunk, 8 unconditional (clamped) loads per thread in flight
      const int ta = rows * m, tb = rows * n;
      if (a.dense4) {
        // both operands dense (lda == m, ldb == n) and 16-byte aligned: the chunk is two contiguous runs,
        // copied as float4 (32 KiB per round trip per CTA instead of 8)
        const float4* ga = reinterpret_cast<const float4*>(g.A + kc * g.lda);
        const float4* gb = reinterpret_cast<const float4*>(g.B + kc * g.ldb);
        float4* sa4 = reinterpret_cast<float4*>(As);
        float4* sb4 = reinterpret_cast<float4*>(Bs);
        const int na = ta >> 2, nb = tb >> 2;
        for (int i0 = t; i0 < na; i0 += 256 * 4) {
          float4 v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) v[u] = __ldg(ga + min(i0 + u * 256, na - 1));
#pragma unroll
          for (int u = 0; u < 4; ++u) { const int i = i0 + u * 256; if (i < na) sa4[i] = v[u]; }
        }
        for (int i0 = t; i0 < nb; i0 += 256 * 8) {
          float4 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = __ldg(gb + min(i0 + u * 256, nb - 1));
#pragma unroll
          for (int u = 0; u < 8; ++u) { const int i = i0 + u * 256; if (i < nb) sb4[i] = v[u]; }
        }
      } else {
      for (int i0 = t; i0 < ta; i0 += 256 * 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int i = min(i0 + u * 256, ta - 1);
          const int rr = i / m, cc = i - rr * m;
          v[u] = __ldg(g.A + (kc + rr) * g.lda + cc);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) { const int i = i0 + u * 256; if (i < ta) As[i] = v[u]; }
      }
      for (int i0 = t; i0 < tb; i0 += 256 * 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int i = min(i0 + u * 256, tb - 1);
          const int rr = i / n, cc = i - rr * n;
          v[u] = __ldg(g.B + (kc + rr) * g.ldb + cc);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) { const int i = i0 + u * 256; if (i < tb) Bs[i] = v[u]; }
      }
      }
    }
    __syncthreads();
    if (active) {
      for (int rr = grp; rr < rows; rr += G) {
        const float* ap = As + rr * m;
        const float* bp = Bs + rr * n;
#pragma unroll
        for (int q = 0; q < Q; ++q) acc[q] = fmaf(ap[ai[q]], bp[bj[q]], acc[q]);
      }
    }
  }
  // groups meet (fixed order), then the CTA's partial goes out
  if (G > 1) {
    __syncthreads();
    red[t] = active ? acc[0] : 0.f;
    __syncthreads();
    if (t < mn) { float r = 0.f; for (int q = 0; q < G; ++q) r += red[q * mn + t]; acc[0] = r; }
  }
  const bool owner = (G > 1) ? t < mn : true;
  if (a.ctas == 1) {
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int o = o0 + 256 * q;
      if (owner && o < mn) g.C[(int64_t)ai[q] * g.ldc + bj[q]] = epilogue(g, acc[q], ai[q], bj[q]);
    }
    return;
  }
  float* mine = a.partial + (int64_t)blockIdx.x * mn;
#pragma unroll
  for (int q = 0; q < Q; ++q) { const int o = o0 + 256 * q; if (owner && o < mn) mine[o] = acc[q]; }
  __syncthreads();
  if (t == 0) {
    __threadfence();
    is_last = (atomicAdd(a.counter, 1u) == (unsigned)a.ctas - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (G > 1) {
    // thread (grp, o) sums partials grp, grp+G, ...; 4 in flight; groups meet in shared memory
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (active) {
      int c = grp;
      for (; c + 3 * G < a.ctas; c += 4 * G) {
        s0 += __ldcg(a.partial + (int64_t)c * mn + o0); s1 += __ldcg(a.partial + (int64_t)(c + G) * mn + o0);
        s2 += __ldcg(a.partial + (int64_t)(c + 2 * G) * mn + o0); s3 += __ldcg(a.partial + (int64_t)(c + 3 * G) * mn + o0);
      }
      for (; c < a.ctas; c += G) s0 += __ldcg(a.partial + (int64_t)c * mn + o0);
    }
    __syncthreads();
    red[t] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (t < mn) {
      float r = 0.f;
      for (int q = 0; q < G; ++q) r += red[q * mn + t];
      g.C[(int64_t)ai[0] * g.ldc + bj[0]] = epilogue(g, r, ai[0], bj[0]);
    }
  } else {
    // Q outputs x 4 partials per step, every load unconditional (clamped) so that all of them are in flight
    float s[Q][4];
#pragma unroll
    for (int q = 0; q < Q; ++q) { s[q][0] = s[q][1] = s[q][2] = s[q][3] = 0.f; }
    for (int c = 0; c < a.ctas; c += 4) {
      float v[Q][4];
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int o = min(o0 + 256 * q, mn - 1);
#pragma unroll
        for (int u = 0; u < 4; ++u) v[q][u] = __ldcg(a.partial + (int64_t)min(c + u, a.ctas - 1) * mn + o);
      }
#pragma unroll
      for (int q = 0; q < Q; ++q)
#pragma unroll
        for (int u = 0; u < 4; ++u) s[q][u] += (c + u < a.ctas) ? v[q][u] : 0.f;
    }
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int o = o0 + 256 * q;
      if (o < mn) g.C[(int64_t)ai[q] * g.ldc + bj[q]] = epilogue(g, (s[q][0] + s[q][1]) + (s[q][2] + s[q][3]), ai[q], bj[q]);
    }
  }
  if (t == 0) *a.counter = 0;
}

static int sgemm_skinny_tn_try(const GemmArgs& g, int64_t batch) {
  if (batch != 1 || !g.ta || g.tb) return 0;
  const int64_t mn = g.m * g.n;
  if (mn > 2048 || g.m + g.n > 176 || g.k < 2048) return 0;        // chunk of A and B rows <= 45 KiB of shared memory
  SkinnyTnArgs a;
  a.g = g;
  // slices: at least 64 rows each; the fold reads ctas x mn floats in one CTA, so fewer slices for bigger results
  int64_t ctas = g.k / 64;
  const bool longk = g.k >= 32768;
  const int64_t cap = mn <= 256 ? (int64_t)S().sm_count * (longk ? 2 : 1) : (mn <= 1024 ? (longk ? 148 : 96) : (longk ? 128 : 64));
  if (ctas > cap) ctas = cap;
  if (ctas < 1) ctas = 1;
  a.rows_per_cta = (g.k + ctas - 1) / ctas;
  ctas = (g.k + a.rows_per_cta - 1) / a.rows_per_cta;
  a.ctas = (int)ctas;
  a.partial = nullptr; a.counter = nullptr;
  if (ctas > 1) {
    const int rc = ensure_scratch((size_t)ctas * mn * sizeof(float));
    if (rc) return -rc;
    a.partial = (float*)scratch_base();
    a.counter = counters_base();
  }
  int ks = (int)(8192 / (g.m + g.n)) & ~7;
  if (ks > 512) ks = 512;
  if (ks < 64) ks = 64;
  a.ks = ks;
  // float4 staging needs dense operands with 16-byte aligned bases and every chunk to start and end on a multiple
  // of 4 rows (ks is a multiple of 8; slices are rounded up to 4 rows; k itself a multiple of 4)
  a.dense4 = 0;
  if (g.lda == g.m && g.ldb == g.n && ((uintptr_t)g.A & 15) == 0 && ((uintptr_t)g.B & 15) == 0 && g.k % 4 == 0) {
    a.rows_per_cta = (a.rows_per_cta + 3) & ~int64_t(3);
    ctas = (g.k + a.rows_per_cta - 1) / a.rows_per_cta;        // never more than the scratch was sized for
    a.ctas = (int)ctas;
    a.dense4 = 1;
  }
  const size_t smem = (size_t)(ks * (g.m + g.n) + 256) * sizeof(float);
  cudaStream_t st = S().stream;
  const int q = (int)((mn + 255) / 256);
  if (q <= 1) sgemm_skinny_tn_kernel<1><<<(unsigned)ctas, 256, smem, st>>>(a);
  else if (q <= 2) sgemm_skinny_tn_kernel<2><<<(unsigned)ctas, 256, smem, st>>>(a);
  else if (q <= 4) sgemm_skinny_tn_kernel<4><<<(unsigned)ctas, 256, smem, st>>>(a);
  else sgemm_skinny_tn_kernel<8><<<(unsigned)ctas, 256, smem, st>>>(a);
  const int rc = post_launch("wf2cu_gemm(skinny tn)");
  return rc ? -rc : 1;
}

}  // namespace wf2

using namespace wf2;

static bool raw_auto() {   // WF2CU_GEMM_RAW=0: auto dispatch keeps the planes kernels (A/B measurements)
  static int on = -1;
  if (on < 0) { const char* e = getenv("WF2CU_GEMM_RAW"); on = (e && e[0] == '0') ? 0 : 1; }
  return on != 0;
}

extern "C" {

int wf2cu_gemm_last_path(int* out) {
  WF2_CHECK(out, "wf2cu_gemm_last_path: null out");
  *out = S().gemm_last_path;
  return 0;
}

int wf2cu_gemm_set_correction(int kind) {
  WF2_CHECK(kind == WF2CU_GEMM_CORRECTION_BF16 || kind == WF2CU_GEMM_CORRECTION_TF32, "wf2cu_gemm_set_correction: kind %d not in {0, 1}", kind);
  gemm_tc_set_correction(kind);
  return 0;
}

int wf2cu_emit_operand_planes(int on) {
  WF2_REQUIRE_INIT();
  S().emit_next = on ? 1 : 0;
  return 0;
}

int wf2cu_gemm_set_mode(int mode) {
  WF2_CHECK(mode >= 0 && mode <= 8, "wf2cu_gemm_set_mode: mode %d not in 0..8", mode);
  S().gemm_mode = mode;
  return 0;
}

static int gemm_impl(int trans_a, int trans_b, int64_t m, int64_t n, int64_t k, const void* A, int64_t a_off, int64_t lda,
                     const void* B, int64_t b_off, int64_t ldb, void* C, int64_t c_off, int64_t ldc, int64_t batch,
                     int64_t stride_a, int64_t stride_b, int64_t stride_c, const float* bias, int act, const float* mask,
                     int64_t ldmask, const char* who, const float* res = nullptr, int64_t ldres = 0) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(A && B && C, "%s: null operand", who);
  WF2_CHECK(m >= 0 && n >= 0 && k >= 0 && batch >= 0, "%s: negative extent", who);
  WF2_CHECK(lda >= (trans_a ? m : k) && ldb >= (trans_b ? k : n) && ldc >= n,
            "%s: leading dimension too small (lda=%lld ldb=%lld ldc=%lld for m=%lld n=%lld k=%lld ta=%d tb=%d)", who,
            (long long)lda, (long long)ldb, (long long)ldc, (long long)m, (long long)n, (long long)k, trans_a, trans_b);
  if (m == 0 || n == 0 || batch == 0) return 0;
  const float* a = (const float*)A + a_off;
  const float* b = (const float*)B + b_off;
  float* c = (float*)C + c_off;
  const bool epi = bias || act || mask;
  WF2_CHECK(!epi || batch == 1, "%s: the fused epilogue takes unbatched operands (batch=%lld)", who, (long long)batch);
  GemmArgs g{a, b, c, m, n, k, lda, ldb, ldc, stride_a, stride_b, stride_c, trans_a ? 1 : 0, trans_b ? 1 : 0, bias, mask, ldmask, act};
  g.res = res; g.ldres = ldres;
  const int mode = S().gemm_mode;
  // gelu / residual epilogues ride in the persistent tensor-core kernels (planes and raw); every other path computes
  // C = A.B + bias first and applies the rest in one element-wise pass (same operations, same order)
  const bool ext = res != nullptr || act == 2;
  auto tail_pass = [&]() -> int {
    GemmArgs e = g;
    e.bias = nullptr;
    int64_t grid = (m * n + 255) / 256;
    const int64_t cap = (int64_t)S().sm_count * 8;
    if (grid > cap) grid = cap;
    gemm_epilogue_kernel<<<(unsigned)grid, 256, 0, S().stream>>>(e);
    return post_launch("wf2cu_gemm_ex2(epilogue)");
  };
  // auto: 50 MFLOP..2.5 GFLOP -> fused tcgen05 kernel (38.5 vs 44.1 us on 1024^3, 89.5 vs 55.6 us on 1280^3; 14.6 vs 22.3 us on 1024x256x784, 8.6 vs 10.2 us on 1024x256x256;
  // below 50 MFLOP its ~5 us floor loses to the SIMT split-K kernel: 7.9 vs 6.9 us on 1024x10x256; gpurun_out/gemm_small5.log)
  // mode: 0 auto | 1 fp32 SIMT | 2 tcgen05 3xTF32 (planes + persistent kernel) | 3 tcgen05 1xTF32 | 4 tcgen05 fused small kernel
  //       | 5 / 6 tcgen05 3xTF32 planes + the 256x256 / 256x192-tile persistent kernel
  if (mode == 0 || mode == 1) {
    // a long reduction with a small result (Conv2D's weight gradient: 16 x 100 over k = 82 944 at batch 1024) goes to
    // the K-sliced kernel whatever its flop count: on a tile-per-CTA kernel it is two tiles (604 us on 16 CTAs)
    const int r = sgemm_skinny_tn_try(g, batch);
    if (r < 0) return -r;
    if (r == 1) { S().gemm_last_path = WF2CU_GEMM_PATH_SIMT_SKINNY; return 0; }
  }
  if (ext && mode != 7 && mode != 2 && mode != 5 && mode != 6 && mode != 8 &&
      !(mode == 0 && 2.0 * (double)m * (double)n * (double)k * (double)batch >= 2.5e9 && m >= 64 && n >= 64 && k >= 32)) {
    // small problems: the plain product (+ bias) on whatever kernel auto picks, then the rest of the epilogue as one pass
    const int rc = gemm_impl(trans_a, trans_b, m, n, k, A, a_off, lda, B, b_off, ldb, C, c_off, ldc, batch, stride_a, stride_b, stride_c, bias, 0,
                             nullptr, 0, who);
    if (rc) return rc;
    return tail_pass();
  }
  {
    const double flops = 2.0 * (double)m * (double)n * (double)k;
    // batched with a narrow result (GPT-2's probabilities x V, n = 64 per head): the 128x256 tile of the planes path
    // wastes 3/4 of every MMA and its split pass rewrites the whole (1024 1024) operand per head twice; the fused
    // kernel converts in flight (414 us -> see profiles/r01_gpt2_launches.md)
    const bool narrow_batched = batch > 1 && n <= 128 && m >= 256 && k >= 256;
    if (mode == 4 || (mode == 0 && ((batch == 1 && flops >= 5.0e7 && flops < 2.5e9) || narrow_batched))) {
      const int r = gemm_tc_fused_try(trans_a ? 1 : 0, trans_b ? 1 : 0, m, n, k, a, lda, b, ldb, c, ldc, batch, stride_a, stride_b,
                                      stride_c, bias, act, mask, ldmask);
      if (r < 0) return -r;
      if (r == 1) return 0;
    }
    if (mode == 4) return set_error(WF2CU_ERR_UNSUPPORTED, "%s: fused tensor-core kernel forced but declined (out of range)", who);
  }
  if (mode == 7 || (mode == 0 && raw_auto() && !trans_a && 2.0 * (double)m * (double)n * (double)k * (double)batch >= 1.0e9 && m >= 64 && n >= 64 &&
                    k >= 32 && gemm_raw_preferred(trans_b ? 1 : 0, m, n, k, batch,
                                                  gemm_tc_prepass_share(a, lda, stride_a, (stride_a == 0 || batch == 1) ? 1 : batch, true),
                                                  gemm_tc_prepass_share(b, ldb, stride_b, (stride_b == 0 || batch == 1) ? 1 : batch, trans_b != 0)))) {
    // A is K-major: raw fp32 tiles by TMA, lo planes made on the SM — no operand pre-pass (gemm_raw.cuh)
    const int r = gemm_raw_try(trans_a, trans_b, m, n, k, a, lda, b, ldb, c, ldc, batch, stride_a, stride_b, stride_c, bias, act, res, ldres);
    if (r < 0) return -r;
    if (r == 1) {
      if (!mask) return 0;
      g.bias = nullptr; g.act = 0; g.res = nullptr;
      int64_t grid = (m * n + 255) / 256;
      const int64_t cap = (int64_t)S().sm_count * 8;
      if (grid > cap) grid = cap;
      gemm_epilogue_kernel<<<(unsigned)grid, 256, 0, S().stream>>>(g);
      return post_launch("wf2cu_gemm_ex(epilogue)");
    }
    if (mode == 7) return set_error(WF2CU_ERR_UNSUPPORTED, "%s: raw-operand tensor-core kernel forced but declined (A transposed or misaligned)", who);
  }
  if (mode != 1) {
    // bias and relu ride in the persistent kernel's epilogue; only the relu-backward mask needs its own pass
    const int r = gemm_tc_try(trans_a, trans_b, m, n, k, a, lda, b, ldb, c, ldc, batch, stride_a, stride_b, stride_c, mode, bias, act, res, ldres);
    if (r < 0) return -r;
    if (r == 1) {
      if (!mask) return 0;
      g.bias = nullptr; g.act = 0; g.res = nullptr;
      int64_t grid = (m * n + 255) / 256;
      const int64_t cap = (int64_t)S().sm_count * 8;
      if (grid > cap) grid = cap;
      gemm_epilogue_kernel<<<(unsigned)grid, 256, 0, S().stream>>>(g);
      return post_launch("wf2cu_gemm_ex(epilogue)");
    }
    if (mode >= 2) return set_error(WF2CU_ERR_UNSUPPORTED, "%s: tensor-core path forced but declined this shape/alignment", who);
  }
  if (ext) {
    const int rc = gemm_impl(trans_a, trans_b, m, n, k, A, a_off, lda, B, b_off, ldb, C, c_off, ldc, batch, stride_a, stride_b, stride_c, bias, 0,
                             nullptr, 0, who);
    if (rc) return rc;
    return tail_pass();
  }
  WF2_CHECK(batch < 65536, "%s: batch %lld too large for the SIMT path", who, (long long)batch);
  {
    int r = sgemm_skinny_nt_try(g, batch);
    if (r == 0) r = sgemm_skinny_tn_try(g, batch);
    if (r < 0) return -r;
    if (r == 1) { S().gemm_last_path = WF2CU_GEMM_PATH_SIMT_SKINNY; return 0; }
  }
  {
    const int r = sgemm_small_try(g, batch);
    if (r < 0) return -r;
    if (r == 1) { S().gemm_last_path = WF2CU_GEMM_PATH_SIMT_SPLITK; return 0; }
  }
  dim3 grid((unsigned)((n + BN - 1) / BN), (unsigned)((m + BM - 1) / BM), (unsigned)batch);
  sgemm_simt_kernel<<<grid, 256, 0, S().stream>>>(g);
  S().gemm_last_path = WF2CU_GEMM_PATH_SIMT;
  return post_launch("wf2cu_gemm(simt)");
}

int wf2cu_gemm_const_operand(const void* base, int64_t offset, int64_t count, int constant) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(base && offset >= 0 && count > 0, "wf2cu_gemm_const_operand: bad range (offset %lld, count %lld)", (long long)offset, (long long)count);
  return gemm_tc_const_declare((const float*)base + offset, (size_t)count * 4, constant);
}

int wf2cu_gemm_const_stats(int64_t* entries, int64_t* bytes) {
  WF2_CHECK(entries && bytes, "wf2cu_gemm_const_stats: null output");
  gemm_tc_const_stats(entries, bytes);
  return 0;
}

int wf2cu_gemm(int trans_a, int trans_b, int64_t m, int64_t n, int64_t k, const void* A, int64_t a_off, int64_t lda,
               const void* B, int64_t b_off, int64_t ldb, void* C, int64_t c_off, int64_t ldc, int64_t batch,
               int64_t stride_a, int64_t stride_b, int64_t stride_c) {
  const int rc_ = gemm_impl(trans_a, trans_b, m, n, k, A, a_off, lda, B, b_off, ldb, C, c_off, ldc, batch, stride_a, stride_b, stride_c,
                   nullptr, 0, nullptr, 0, "wf2cu_gemm");
  S().emit_next = 0;      // wf2cu_emit_operand_planes is one shot: a product that could not write the plane drops the request
  return rc_;
}

int wf2cu_gemm_ex(int trans_a, int trans_b, int64_t m, int64_t n, int64_t k, const void* A, int64_t a_off, int64_t lda,
                  const void* B, int64_t b_off, int64_t ldb, void* C, int64_t c_off, int64_t ldc,
                  const void* bias, int64_t bias_off, int act, const void* mask, int64_t mask_off, int64_t ldmask) {
  WF2_CHECK(act == 0 || act == 1, "wf2cu_gemm_ex: act %d not in {0 none, 1 relu}", act);
  WF2_CHECK(!mask || ldmask >= n, "wf2cu_gemm_ex: ldmask %lld < n %lld", (long long)ldmask, (long long)n);
  const int rc_ = gemm_impl(trans_a, trans_b, m, n, k, A, a_off, lda, B, b_off, ldb, C, c_off, ldc, 1, 0, 0, 0,
                   bias ? (const float*)bias + bias_off : nullptr, act, mask ? (const float*)mask + mask_off : nullptr, ldmask,
                   "wf2cu_gemm_ex");
  S().emit_next = 0;      // wf2cu_emit_operand_planes is one shot: a product that could not write the plane drops the request
  return rc_;
}

int wf2cu_gemm_ex2(int trans_a, int trans_b, int64_t m, int64_t n, int64_t k, const void* A, int64_t a_off, int64_t lda,
                   const void* B, int64_t b_off, int64_t ldb, void* C, int64_t c_off, int64_t ldc,
                   const void* bias, int64_t bias_off, int act, const void* residual, int64_t res_off, int64_t ldres) {
  WF2_CHECK(act >= 0 && act <= 2, "wf2cu_gemm_ex2: act %d not in {0 none, 1 relu, 2 gelu}", act);
  WF2_CHECK(!residual || ldres >= n, "wf2cu_gemm_ex2: ldres %lld < n %lld", (long long)ldres, (long long)n);
  const int rc_ = gemm_impl(trans_a, trans_b, m, n, k, A, a_off, lda, B, b_off, ldb, C, c_off, ldc, 1, 0, 0, 0,
                   bias ? (const float*)bias + bias_off : nullptr, act, nullptr, 0, "wf2cu_gemm_ex2",
                   residual ? (const float*)residual + res_off : nullptr, ldres);
  S().emit_next = 0;      // wf2cu_emit_operand_planes is one shot: a product that could not write the plane drops the request
  return rc_;
}

}  // extern "C"
