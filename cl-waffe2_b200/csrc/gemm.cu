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
#include "common.cuh"

namespace wf2 {

// implemented in gemm_tc.cu; returns 1 when it handled the problem, 0 when it declines, <0 error
int gemm_tc_try(int ta, int tb, int64_t m, int64_t n, int64_t k, const float* A, int64_t lda, const float* B,
                int64_t ldb, float* C, int64_t ldc, int64_t batch, int64_t sa, int64_t sb, int64_t sc, int mode);

constexpr int BM = 128, BN = 128, BK = 16;

struct GemmArgs {
  const float* A; const float* B; float* C;
  int64_t m, n, k, lda, ldb, ldc, sa, sb, sc;
  int ta, tb;
};

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
      if (gn < g.n) C[gm * g.ldc + gn] = acc[i][j];
    }
  }
}

}  // namespace wf2

using namespace wf2;

extern "C" {

int wf2cu_gemm_set_mode(int mode) {
  WF2_CHECK(mode >= 0 && mode <= 3, "wf2cu_gemm_set_mode: mode %d not in 0..3", mode);
  S().gemm_mode = mode;
  return 0;
}

int wf2cu_gemm(int trans_a, int trans_b, int64_t m, int64_t n, int64_t k, const void* A, int64_t a_off, int64_t lda,
               const void* B, int64_t b_off, int64_t ldb, void* C, int64_t c_off, int64_t ldc, int64_t batch,
               int64_t stride_a, int64_t stride_b, int64_t stride_c) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(A && B && C, "wf2cu_gemm: null operand");
  WF2_CHECK(m >= 0 && n >= 0 && k >= 0 && batch >= 0, "wf2cu_gemm: negative extent");
  WF2_CHECK(lda >= (trans_a ? m : k) && ldb >= (trans_b ? k : n) && ldc >= n,
            "wf2cu_gemm: leading dimension too small (lda=%lld ldb=%lld ldc=%lld for m=%lld n=%lld k=%lld ta=%d tb=%d)",
            (long long)lda, (long long)ldb, (long long)ldc, (long long)m, (long long)n, (long long)k, trans_a, trans_b);
  if (m == 0 || n == 0 || batch == 0) return 0;
  const float* a = (const float*)A + a_off;
  const float* b = (const float*)B + b_off;
  float* c = (float*)C + c_off;
  const int mode = S().gemm_mode;
  if (mode != 1) {
    const int r = gemm_tc_try(trans_a, trans_b, m, n, k, a, lda, b, ldb, c, ldc, batch, stride_a, stride_b, stride_c, mode);
    if (r < 0) return -r;
    if (r == 1) return 0;
    if (mode >= 2) return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_gemm: tensor-core path forced but declined this shape/alignment");
  }
  WF2_CHECK(batch < 65536, "wf2cu_gemm: batch %lld too large for the SIMT path", (long long)batch);
  GemmArgs g{a, b, c, m, n, k, lda, ldb, ldc, stride_a, stride_b, stride_c, trans_a ? 1 : 0, trans_b ? 1 : 0};
  dim3 grid((unsigned)((n + BN - 1) / BN), (unsigned)((m + BM - 1) / BM), (unsigned)batch);
  sgemm_simt_kernel<<<grid, 256, 0, S().stream>>>(g);
  return post_launch("wf2cu_gemm(simt)");
}

}  // extern "C"
