// gemm_tc.cu — MatmulNode on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), fp32-faithful.
//
// The reference runs `!matmul` as Fortran `sgemm_` in fp32 (`source/backends/cpu/matrix-ops.lisp:58-79`).
// tcgen05 has no fp32-input MMA and one TF32 pass misses the 1e-4 bar (2.5e-4), so
// this path is the error-compensated 3xTF32 split:
//     A = A_hi + A_lo,  B = B_hi + B_lo   (hi = tf32_rn(x), lo = tf32_rn(x - hi))
//     C = A_hi.B_hi + A_hi.B_lo + A_lo.B_hi        accumulated in fp32 in TMEM
// Two launches per MatmulNode:
//   1. split_planes_kernel  : one pass over each operand writing its hi / lo planes K-major
//      ([batch][rows][Kp], Kp = K rounded up to 32, zero padded).  An operand whose storage is
//      MN-major for the GEMM (A with trans_a, B without trans_b) is transposed here through
//      shared memory, so step 2 sees a single layout whatever the LazyTransposeNode flags were.
//   2. gemm_tc_kernel       : persistent (one CTA per SM), warp-specialised:
//        warp 0    TMA producer   cp.async.bulk.tensor.3d, SWIZZLE_128B, 4 planes per stage
//        warp 1    MMA issuer     tcgen05.mma.cta_group::1.kind::tf32, 128x256x8, 3 per k-slice,
//                                 accumulator double-buffered in TMEM (2 x 256 columns)
//        warps 2-5 epilogue       tcgen05.ld 32x32b.x32 -> registers -> SWIZZLE_128B staging block in shared memory ->
//                                 cp.async.bulk.tensor store through a tensor map of C (clipped at the edges by the map);
//                                 C layouts a tensor map cannot describe are stored from registers, bounds-checked
//      smem full/empty mbarriers between warp 0 and warp 1, TMEM full/empty between 1 and 2-5.
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>
#include <map>
#include <tuple>
#include <type_traits>
#include <vector>
#include "common.cuh"
#include "math.cuh"
#include "tc.cuh"

namespace wf2 {

constexpr int TM = 128;          // tile rows  (UMMA M)
constexpr int TN = 256;          // tile cols  (UMMA N)
constexpr int TK = 32;           // fp32 elements per k-block = one 128-byte swizzle row
constexpr int UK = 8;            // UMMA K for kind::tf32 (32 bytes)
constexpr int STAGES = 2;
constexpr int A_PLANE_BYTES = TM * TK * 4;   // 16 KiB
constexpr int B_PLANE_BYTES = TN * TK * 4;   // 32 KiB
constexpr int STAGE_BYTES = 2 * A_PLANE_BYTES + 2 * B_PLANE_BYTES;   // 96 KiB
constexpr int EPI_LD = 36;                               // floats per staged output row (32 + 4: float4 rows stay conflict-free)
constexpr int EPI_SLOT = 8192;                           // per epilogue warp, 1024-byte aligned: two 4 KiB 32 x 32 blocks in the
                                                         // SWIZZLE_128B layout of the bulk tensor store (or one padded block, 4608 B)
constexpr int EPI_BYTES = 4 * EPI_SLOT;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + EPI_BYTES + 256 /*barriers*/;
static_assert(32 * EPI_LD * 4 <= EPI_SLOT && SMEM_BYTES <= 232448, "shared-memory budget");
constexpr int GEMM_THREADS = 192;

// instruction descriptor (cute::UMMA::InstrDescriptor): D=f32 [4,6)=1, A=tf32 [7,10)=2, B=tf32 [10,13)=2,
// both K-major (bits 15,16 = 0), N>>3 [17,23), M>>4 [24,29)
constexpr uint32_t kIdescTf32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
// the same for kind::f16 with bf16 operands: A = B = bf16 (format 1), K = 16 per instruction
__host__ __device__ constexpr uint32_t idesc_bf16_mn(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
constexpr uint32_t kIdescBf16 = idesc_bf16_mn(TM, TN);

__device__ __forceinline__ float tf32_rn(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ void split1(float x, float& hi, float& lo) {
  hi = tf32_rn(x);
  const float d = x - hi;              // exact
  lo = (fabsf(hi) <= 3.0e38f) ? tf32_rn(d) : 0.f;   // inf/nan live in hi only
}

// ------------------------------------------------------------------ step 1: hi / lo planes
// Two forms of the second plane:
//   bf_blk == 0   lo = tf32_rn(x - hi), fp32: the classic 3xTF32 split (C = hi.hi + hi.lo + lo.hi, three kind::tf32 MMAs)
//   bf_blk == 16 / 32   a MIXED bf16 plane of the same size and geometry: per k-block of bf_blk elements (bf_blk * 4 bytes, one
//                 swizzle row of the consuming kernel) first bf16_rn(x) of the block, then bf16_rn(x - hi) of the block.  The
//                 corrections hi.lo + lo.hi then run as TWO kind::f16 (bf16) MMAs of K = 16 instead of four kind::tf32 MMAs of
//                 K = 8: the tensor pipe spends 2 units per 8 k instead of 3.  The corrections are ~2^-12 of the product, so the
//                 bf16 rounding of their factors (2^-9) leaves ~2^-20 per product: measured 5e-7 of max|C| on random operands,
//                 the size of an fp32 sgemm's own rounding error (scripts/emulate_hybrid_split.py).
struct SplitArgs {
  const float* src;   // operand base (already offset)
  float* hi;          // [batch][R][Kp]
  float* lo;
  int64_t ld, sx;     // leading dimension and batch stride of src (elements)
  int R, K, Kp;       // rows of the plane (m or n), true k, padded k
  int vec;            // src rows are 16-byte aligned (direct variant only)
  int bf_blk;         // 0: fp32 lo plane ; 16 / 32: mixed bf16 plane with that k-block
  int raw_hi;         // the kernel reads the SOURCE itself as its hi plane (TMA map over the operand; kind::tf32 uses the upper
                      // 19 bits of the container, i.e. hi = trunc_tf32(x)): no hi plane is written and lo = x - trunc_tf32(x)
};

__device__ __forceinline__ float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }

template <bool RAW_HI>
__device__ __forceinline__ void split_bf(float x, float& hi, uint16_t& hb, uint16_t& lb) {
  hi = RAW_HI ? trunc_tf32(x) : tf32_rn(x);
  if (fabsf(hi) <= 3.0e38f) {
    hb = __bfloat16_as_ushort(__float2bfloat16_rn(x));
    lb = __bfloat16_as_ushort(__float2bfloat16_rn(x - hi));   // x - hi is exact
  } else {             // inf / nan live in hi only: the corrections must not turn inf into inf - inf or inf . 0
    hb = 0; lb = 0;
  }
}
template <bool RAW_HI>
__device__ __forceinline__ void split1v(float x, float& hi, float& lo) {
  if (!RAW_HI) { split1(x, hi, lo); return; }
  hi = trunc_tf32(x);
  lo = (fabsf(x) <= 3.0e38f) ? tf32_rn(x - hi) : 0.f;
}

// src is K-contiguous: (r, k) at r*ld + k
template <bool RAW_HI>
__global__ void __launch_bounds__(256) split_planes_direct_kernel(const __grid_constant__ SplitArgs a) {
  const int kq = a.Kp >> 2;
  const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (int64_t)a.R * kq) return;
  const int r = (int)(idx / kq), k = (int)(idx % kq) * 4;
  const int64_t b = blockIdx.y;
  const float* s = a.src + b * a.sx + (int64_t)r * a.ld + k;
  float x[4];
  if (a.vec && k + 3 < a.K) {
    const float4 v = *reinterpret_cast<const float4*>(s);
    x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j) x[j] = (k + j < a.K) ? s[j] : 0.f;
  }
  const int64_t o = (b * a.R + r) * (int64_t)a.Kp + k;
  if (a.bf_blk) {
    float4 h;
    uint16_t hb[4], lb[4];
    split_bf<RAW_HI>(x[0], h.x, hb[0], lb[0]); split_bf<RAW_HI>(x[1], h.y, hb[1], lb[1]);
    split_bf<RAW_HI>(x[2], h.z, hb[2], lb[2]); split_bf<RAW_HI>(x[3], h.w, hb[3], lb[3]);
    if (!RAW_HI) *reinterpret_cast<float4*>(a.hi + o) = h;
    // row base in 16-bit units: 2 * Kp per row; k-block kb holds hb[0:blk] then lb[0:blk]
    uint16_t* row = reinterpret_cast<uint16_t*>(a.lo) + (b * a.R + r) * (int64_t)(2 * a.Kp);
    const int kb = k / a.bf_blk, j = k - kb * a.bf_blk;
    uint16_t* dst = row + (int64_t)kb * 2 * a.bf_blk + j;
    *reinterpret_cast<uint2*>(dst) = make_uint2((uint32_t)hb[0] | ((uint32_t)hb[1] << 16), (uint32_t)hb[2] | ((uint32_t)hb[3] << 16));
    *reinterpret_cast<uint2*>(dst + a.bf_blk) = make_uint2((uint32_t)lb[0] | ((uint32_t)lb[1] << 16), (uint32_t)lb[2] | ((uint32_t)lb[3] << 16));
    return;
  }
  float4 h, l;
  split1v<RAW_HI>(x[0], h.x, l.x); split1v<RAW_HI>(x[1], h.y, l.y); split1v<RAW_HI>(x[2], h.z, l.z); split1v<RAW_HI>(x[3], h.w, l.w);
  if (!RAW_HI) *reinterpret_cast<float4*>(a.hi + o) = h;
  *reinterpret_cast<float4*>(a.lo + o) = l;
}

// src is R-contiguous: (r, k) at k*ld + r  -> transpose through shared memory
__global__ void __launch_bounds__(256) split_planes_transpose_kernel(const __grid_constant__ SplitArgs a) {
  __shared__ float th[32][33];
  __shared__ float tl[32][33];          // fp32 lo, or (hb | lb << 16) of the mixed plane
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  const int r0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const int64_t b = blockIdx.z;
  const float* s = a.src + b * a.sx;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int k = k0 + ty + j * 8, r = r0 + tx;
    const float x = (k < a.K && r < a.R) ? s[(int64_t)k * a.ld + r] : 0.f;
    float h, l;
    if (a.bf_blk) {
      uint16_t hb, lb;
      split_bf<false>(x, h, hb, lb);
      l = __uint_as_float((uint32_t)hb | ((uint32_t)lb << 16));
    } else {
      split1(x, h, l);
    }
    th[ty + j * 8][tx] = h;
    tl[ty + j * 8][tx] = l;
  }
  __syncthreads();
  // Store side: every thread owns FOUR consecutive k of one row (8 threads per row, 4 rows per warp): one 16-byte store into the hi
  // plane and one 16-byte (fp32 lo) or two 8-byte (mixed bf16) stores into the second plane, so a warp writes 4 x 128 contiguous
  // bytes per instruction.  (The first version wrote one element per thread: 4-byte stores into hi and 2-byte stores into the
  // mixed plane — the transposing pre-pass of a plain (k n) B operand cost 26 us more than the direct one at 4096^2.)
  // Shared-memory reads th[4 kq + i][r]: bank (4 kq + i + r) mod 32 over kq = lane % 8, r = 4 w + lane / 8 is a permutation.
  {
    const int rl = threadIdx.x >> 3, kq = threadIdx.x & 7;
    const int r = r0 + rl, k = k0 + 4 * kq;
    if (r < a.R) {   // k < Kp always (Kp is a multiple of 32)
      const int64_t o = (b * a.R + r) * (int64_t)a.Kp + k;
      *reinterpret_cast<float4*>(a.hi + o) = make_float4(th[4 * kq][rl], th[4 * kq + 1][rl], th[4 * kq + 2][rl], th[4 * kq + 3][rl]);
      if (a.bf_blk) {
        const uint32_t w0 = __float_as_uint(tl[4 * kq][rl]), w1 = __float_as_uint(tl[4 * kq + 1][rl]);
        const uint32_t w2 = __float_as_uint(tl[4 * kq + 2][rl]), w3 = __float_as_uint(tl[4 * kq + 3][rl]);
        uint16_t* row = reinterpret_cast<uint16_t*>(a.lo) + (b * a.R + r) * (int64_t)(2 * a.Kp);
        const int kb = k / a.bf_blk, jj = k - kb * a.bf_blk;
        uint16_t* dst = row + (int64_t)kb * 2 * a.bf_blk + jj;
        *reinterpret_cast<uint2*>(dst) = make_uint2((w0 & 0xFFFFu) | (w1 << 16), (w2 & 0xFFFFu) | (w3 << 16));
        *reinterpret_cast<uint2*>(dst + a.bf_blk) = make_uint2((w0 >> 16) | (w1 & 0xFFFF0000u), (w2 >> 16) | (w3 & 0xFFFF0000u));
      } else {
        *reinterpret_cast<float4*>(a.lo + o) = make_float4(tl[4 * kq][rl], tl[4 * kq + 1][rl], tl[4 * kq + 2][rl], tl[4 * kq + 3][rl]);
      }
    }
  }
}

// ------------------------------------------------------------------ step 2: the tcgen05 GEMM
struct TcArgs {
  float* C;
  int64_t ldc, sc;       // C leading dimension and batch stride (elements)
  int m, n;
  int nkb;               // Kp / TK
  int tiles_m, tiles_n, batch;
  int passes;            // 3 = 3xTF32 (fp32 lo planes), 2 = TF32 + two bf16 correction MMAs (mixed bf16 planes; the default),
                         // 1 = single TF32 pass (hi planes only, not parity grade)
  int c_vec;             // C rows are 16-byte aligned -> 128-bit stores
  int a_bcast, b_bcast;  // operand shared by every batch index (batch stride 0): one plane set
  const float* bias;     // != NULL: C[r, c] = acc + bias[c]  (LinearLayer's bias add folded into the epilogue)
  int act;               // 1: then C = C * (C > 0 ? 1 : 0)  (`!relu`) ; 2: C = gelu(C) (`!gelu`, tanh form)
  const float* res;      // != NULL: then C = res[r, c] + C  (a residual connection riding on the projection)
  int64_t ldres;
  int c_tma;             // C goes out through map_c (bulk tensor stores) instead of per-lane global stores
  int group_m = 0;       // m-tiles per raster group (0 = kGroupM)
  int plane = 0;         // 1: the epilogue also writes the mixed bf16 operand plane of C through pmap (wf2cu_emit_operand_planes; 256-row kernels)
  alignas(64) CUtensorMap pmap;   // the plane as an (m x n) fp32-sized tensor, same boxes as map_c
};

// epilogue of the persistent kernels: 32 consecutive columns of one row per lane; same arithmetic as the separate
// nodes.  Called by the whole warp (col0 is warp-uniform): lane j fetches bias[col0 + j] with one coalesced load
// and the 32 values travel by shuffle — 32 dependent loads per lane made the epilogue longer than a K = 768 mainloop.
__device__ __forceinline__ void tc_epilogue32(const TcArgs& a, uint32_t (&v)[32], int row0, int col0, int lane) {
  if (a.bias) {
    const float mine = (col0 + lane < a.n) ? __ldg(a.bias + col0 + lane) : 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      v[j] = __float_as_uint(__fadd_rn(__uint_as_float(v[j]), __shfl_sync(0xffffffffu, mine, j)));
  }
  if (a.act == 1) {
#pragma unroll
    for (int j = 0; j < 32; ++j) { const float x = __uint_as_float(v[j]); v[j] = __float_as_uint(__fmul_rn(x, x > 0.f ? 1.f : 0.f)); }
  } else if (a.act == 2) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(gelu_tanh_ref(__uint_as_float(v[j])));
  }
  if (a.res && row0 + lane < a.m) {          // lane = row: 32 consecutive floats of the residual's row (whole 128-byte lines)
    const float* r = a.res + (int64_t)(row0 + lane) * a.ldres + col0;
    if (col0 + 32 <= a.n && (a.ldres & 7) == 0 && (reinterpret_cast<uintptr_t>(a.res) & 31) == 0 && (col0 & 7) == 0) {
      // 256-bit loads: a load instruction costs one L1 tag per row it touches (32 here), so four instead of eight per 32 columns
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float x[8];
        asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=f"(x[0]), "=f"(x[1]), "=f"(x[2]), "=f"(x[3]), "=f"(x[4]), "=f"(x[5]), "=f"(x[6]), "=f"(x[7]) : "l"(r + 8 * j));
#pragma unroll
        for (int i = 0; i < 8; ++i) v[8 * j + i] = __float_as_uint(__fadd_rn(x[i], __uint_as_float(v[8 * j + i])));
      }
    } else if (col0 + 32 <= a.n && (a.ldres & 3) == 0 && (reinterpret_cast<uintptr_t>(a.res) & 15) == 0 && (col0 & 3) == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 x = __ldg(reinterpret_cast<const float4*>(r) + j);
        v[4 * j] = __float_as_uint(__fadd_rn(x.x, __uint_as_float(v[4 * j])));
        v[4 * j + 1] = __float_as_uint(__fadd_rn(x.y, __uint_as_float(v[4 * j + 1])));
        v[4 * j + 2] = __float_as_uint(__fadd_rn(x.z, __uint_as_float(v[4 * j + 2])));
        v[4 * j + 3] = __float_as_uint(__fadd_rn(x.w, __uint_as_float(v[4 * j + 3])));
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < a.n) v[j] = __float_as_uint(__fadd_rn(__ldg(r + j), __uint_as_float(v[j])));
    }
  }
}

// 32 consecutive columns of row (row0 + lane), held one row per lane -> C.  A lane-per-row float4 store touches 32
// lines with 16 bytes each (32 LSU wavefronts per instruction: ~16K for a 256x256 tile, which is what made the
// K = 64 products of attention and the non-overlapped epilogue of the 256-row tiles slow).  The block is transposed
// through shared memory so that every store instruction writes four whole 128-byte lines.
__device__ __forceinline__ void tc_store32(const TcArgs& a, float* cbase, float* stage, const uint32_t (&v)[32], int row0,
                                           int col0, int lane) {
  if (col0 >= a.n) return;                                  // warp-uniform
  if (a.c_vec && col0 + 32 <= a.n) {
    float* mine = stage + lane * EPI_LD;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<float4*>(mine + 4 * j) =
          make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
    __syncwarp();
    const int seg = lane & 7, sub = lane >> 3;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = 4 * i + sub;
      const float4 x = *reinterpret_cast<const float4*>(stage + r * EPI_LD + 4 * seg);
      if (row0 + r < a.m) *reinterpret_cast<float4*>(cbase + (int64_t)(row0 + r) * a.ldc + col0 + 4 * seg) = x;
    }
    __syncwarp();
  } else if (row0 + lane < a.m) {
    float* crow = cbase + (int64_t)(row0 + lane) * a.ldc;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (col0 + j < a.n) crow[col0 + j] = __uint_as_float(v[j]);
  }
}

// The same block through the bulk-copy engine: each lane writes its row into a 4 KiB staging block in the SWIZZLE_128B
// layout (16-byte chunk j of row r at chunk j ^ (r & 7): conflict-free for the quarter-warps of a 128-bit store), one
// lane hands the block to cp.async.bulk.tensor, which writes whole lines and clips rows >= m / columns >= n.  Per block
// the warp issues 8 shared stores instead of 8 shared stores + 8 shared loads + 8 global stores, and never waits for
// the global stores: two staging blocks alternate, a block is reused once the engine has read it (wait_group.read 1).
__device__ __forceinline__ void tc_store32_tma(const TcArgs& a, const CUtensorMap* map_c, uint8_t* slot, uint32_t& which,
                                               const uint32_t (&v)[32], int row0, int col0, int b, int lane) {
  if (col0 >= a.n || row0 >= a.m) return;                   // warp-uniform
  uint8_t* blk = slot + which * 4096;
  which ^= 1;
  if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
  __syncwarp();
  uint8_t* mine = blk + lane * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    *reinterpret_cast<float4*>(mine + ((j ^ (lane & 7)) << 4)) =
        make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  if (lane == 0) tma_store_3d(map_c, blk, col0, row0, b);
}

// The 256-row kernels (gemm_tc256_kernel, gemm_tc2sm_kernel) keep two 256-column accumulators — all of tensor memory — so their
// epilogue is NOT overlapped with the next tile's main loop; at K = 768 (GPT-2's projections) four epilogue warps made it a third of
// the tile.  They run EIGHT: two per TMEM lane quarter, half of the tile's columns each, one 4 KiB staging block per warp.
constexpr int EPI8_WARPS = 8, EPI8_THREADS = 64 + 32 * EPI8_WARPS, EPI8_SLOT = EPI_BYTES / EPI8_WARPS;
// one staging block per warp: the block is handed to the bulk-copy engine and reused once the engine has read it (the next
// 32 columns' bias / activation arithmetic runs in between)
__device__ __forceinline__ void tc_store32_tma1(const TcArgs& a, const CUtensorMap* map_c, uint8_t* blk, const uint32_t (&v)[32], int row0,
                                                int col0, int b, int lane) {
  if (col0 >= a.n || row0 >= a.m) return;                   // warp-uniform
  if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  __syncwarp();
  uint8_t* mine = blk + lane * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    *reinterpret_cast<float4*>(mine + ((j ^ (lane & 7)) << 4)) =
        make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  if (lane == 0) tma_store_3d(map_c, blk, col0, row0, b);
}
// C cannot go through a tensor map (misaligned base or pitch): per-lane stores of the lane's own row
__device__ __forceinline__ void tc_store32_rows(const TcArgs& a, float* cbase, const uint32_t (&v)[32], int row0, int col0, int lane) {
  if (col0 >= a.n || row0 + lane >= a.m) return;
  float* crow = cbase + (int64_t)(row0 + lane) * a.ldc;
#pragma unroll
  for (int j = 0; j < 32; ++j)
    if (col0 + j < a.n) crow[col0 + j] = __uint_as_float(v[j]);
}

// C is the K-major left operand of the MatmulNode issued next (GPT-2: gelu(fc h) feeds the second MLP projection): the lane's 32
// finished values also leave as the mixed bf16 plane that product's pre-pass would write (plane_bits == split_bf<true>, bit for bit),
// so that pre-pass — a 12-bytes-per-element launch — is not run.  32 columns = two k-blocks of 16 = 128 bytes of plane per row
// ([16 bf16(x)][16 bf16(x - trunc x)] twice): the same staging block and bulk tensor store as C, through the plane's own map.
// (Per-lane 8-byte global stores from here cost 93 us per GPT-2 layer against the 38 us pre-pass they replace: 32 L1 tags per
// instruction.  With four epilogue warps even this form lost, 9.64 -> 9.76 ms per forward; with eight it wins.)
__device__ __forceinline__ void tc_plane32(const TcArgs& a, uint8_t* blk, const uint32_t (&v)[32], int row0, int col0, int b, int lane) {
  if (!a.plane) return;
  uint32_t w[32];
#pragma unroll
  for (int c = 0; c < 32; c += 2) {
    uint32_t h0, l0, h1, l1;
    plane_bits(__uint_as_float(v[c]), h0, l0);
    plane_bits(__uint_as_float(v[c + 1]), h1, l1);
    w[(c >> 4) * 16 + ((c & 15) >> 1)] = h0 | (h1 << 16);
    w[(c >> 4) * 16 + 8 + ((c & 15) >> 1)] = l0 | (l1 << 16);
  }
  tc_store32_tma1(a, &a.pmap, blk, w, row0, col0, b, lane);
}

constexpr int kGroupM = 16;   // raster: 16 m-tiles x all n-tiles per group, m fastest (L2 reuse of B panels)

__device__ __forceinline__ void tile_coords(int t, const TcArgs& a, int& b, int& mt, int& nt) {
  const int per_batch = a.tiles_m * a.tiles_n;
  b = t / per_batch;
  int r = t - b * per_batch;
  const int G = a.group_m > 0 ? a.group_m : kGroupM;
  const int group_sz = G * a.tiles_n;
  const int g = r / group_sz;
  const int first_m = g * G;
  const int gm = min(G, a.tiles_m - first_m);
  r -= g * group_sz;
  mt = first_m + r % gm;
  nt = r / gm;
}

__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap map_ah, const __grid_constant__ CUtensorMap map_al,
               const __grid_constant__ CUtensorMap map_bh, const __grid_constant__ CUtensorMap map_bl,
               const __grid_constant__ CUtensorMap map_c, const __grid_constant__ TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operands need 1024-byte aligned tiles
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-byte aligned, and still known to be SHARED (LDS / STS, not generic LD / ST)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + EPI_BYTES);
  uint64_t* full = bars;                    // [STAGES]  TMA -> MMA
  uint64_t* empty = bars + STAGES;          // [STAGES]  MMA -> TMA
  uint64_t* acc_full = bars + 2 * STAGES;   // [2]       MMA -> epilogue
  uint64_t* acc_empty = acc_full + 2;       // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = a.tiles_m * a.tiles_n * a.batch;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {   // TMEM: all 512 columns (this kernel owns the SM: ~193 KB of shared memory)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      const uint32_t tx_bytes = a.passes != 1 ? STAGE_BYTES : (A_PLANE_BYTES + B_PLANE_BYTES);
      int s = 0; uint32_t ph = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        int b, mt, nt;
        tile_coords(t, a, b, mt, nt);
        for (int kb = 0; kb < a.nkb; ++kb) {
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* st = smem + s * STAGE_BYTES;
          mbar_expect_tx(&full[s], tx_bytes);
          const int ba = a.a_bcast ? 0 : b, bb = a.b_bcast ? 0 : b;
          tma_load_3d(&map_ah, &full[s], st, kb * TK, mt * TM, ba);
          tma_load_3d(&map_bh, &full[s], st + 2 * A_PLANE_BYTES, kb * TK, nt * TN, bb);
          if (a.passes != 1) {
            tma_load_3d(&map_al, &full[s], st + A_PLANE_BYTES, kb * TK, mt * TM, ba);
            tma_load_3d(&map_bl, &full[s], st + 2 * A_PLANE_BYTES + B_PLANE_BYTES, kb * TK, nt * TN, bb);
          }
          if (++s == STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (elect_one()) {
      int s = 0; uint32_t ph = 0;
      int as = 0; uint32_t aph = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        mbar_wait(&acc_empty[as], aph ^ 1);     // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)(as * TN);
        for (int kb = 0; kb < a.nkb; ++kb) {
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t st = smem_u32(smem + s * STAGE_BYTES);
          const uint64_t dah = make_desc_sw128(st), dal = make_desc_sw128(st + A_PLANE_BYTES);
          const uint64_t dbh = make_desc_sw128(st + 2 * A_PLANE_BYTES), dbl = make_desc_sw128(st + 2 * A_PLANE_BYTES + B_PLANE_BYTES);
          if (a.passes == 2) {
            // per 16 k: the two corrections as bf16 MMAs of K = 16 (bf16(A) . bf16(B_lo) + bf16(A_lo) . bf16(B)), then the main
            // product as two TF32 MMAs of K = 8 — the order gemm_tc256_kernel uses per k-block of 16, so that every tile shape
            // accumulates in the same order.  The mixed plane is laid out per k-block of 16 for every kernel (hb[0:16] lb[0:16]
            // hb[16:32] lb[16:32] in this kernel's 128-byte swizzle row, 32 bytes = two 16-byte descriptor units each), so one
            // plane set serves whichever tile shape consumes it (and a producer can write it without knowing the consumer).
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              tc_mma_bf16(d, dal + (uint64_t)(4 * j), dbl + (uint64_t)(4 * j + 2), kIdescBf16, (kb | j) != 0);
              tc_mma_bf16(d, dal + (uint64_t)(4 * j + 2), dbl + (uint64_t)(4 * j), kIdescBf16, 1);
              tc_mma_tf32(d, dah + (uint64_t)(4 * j), dbh + (uint64_t)(4 * j), kIdescTf32, 1);
              tc_mma_tf32(d, dah + (uint64_t)(4 * j + 2), dbh + (uint64_t)(4 * j + 2), kIdescTf32, 1);
            }
          } else {
#pragma unroll
            for (int k = 0; k < TK / UK; ++k) {
              const uint64_t adv = (uint64_t)((k * UK * 4) >> 4);   // +32 bytes per UMMA_K inside the swizzle row
              if (a.passes == 3) {   // small terms first, then the main product
                tc_mma_tf32(d, dal + adv, dbh + adv, kIdescTf32, (kb | k) != 0);
                tc_mma_tf32(d, dah + adv, dbl + adv, kIdescTf32, 1);
                tc_mma_tf32(d, dah + adv, dbh + adv, kIdescTf32, 1);
              } else {
                tc_mma_tf32(d, dah + adv, dbh + adv, kIdescTf32, (kb | k) != 0);
              }
            }
          }
          tc_commit(&empty[s]);                 // frees the smem stage when these MMAs retire
          if (++s == STAGES) { s = 0; ph ^= 1; }
        }
        tc_commit(&acc_full[as]);               // accumulator complete -> epilogue
        if (++as == 2) { as = 0; aph ^= 1; }
      }
    }
  } else {
    // ===================== epilogue: TMEM -> registers -> global =====================
    const int q = warp & 3;                     // TMEM lane quarter this warp may read
    uint8_t* slot = smem + STAGES * STAGE_BYTES + (warp - 2) * EPI_SLOT;
    float* stage = reinterpret_cast<float*>(slot);
    uint32_t which = 0;
    int as = 0; uint32_t aph = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      int b, mt, nt;
      tile_coords(t, a, b, mt, nt);
      mbar_wait(&acc_full[as], aph);
      tc_fence_after();
      float* cbase = a.C + (int64_t)b * a.sc;
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * TN);
      // two register buffers: the TMEM read of the next 32 columns is in flight while this one is stored
      uint32_t v0[32], v1[32];
      tc_ld32(tbase, v0);
#pragma unroll 1
      for (int c = 0; c < TN / 32; c += 2) {
        tc_ld_wait();
        tc_ld32(tbase + (c + 1) * 32, v1);
        int col0 = nt * TN + c * 32;
        if (col0 < a.n) tc_epilogue32(a, v0, mt * TM + q * 32, col0, lane);
        if (a.c_tma) tc_store32_tma(a, &map_c, slot, which, v0, mt * TM + q * 32, col0, b, lane);
        else tc_store32(a, cbase, stage, v0, mt * TM + q * 32, col0, lane);
        tc_ld_wait();
        if (c + 2 < TN / 32) tc_ld32(tbase + (c + 2) * 32, v0);
        col0 += 32;
        if (col0 < a.n) tc_epilogue32(a, v1, mt * TM + q * 32, col0, lane);
        if (a.c_tma) tc_store32_tma(a, &map_c, slot, which, v1, mt * TM + q * 32, col0, b, lane);
        else tc_store32(a, cbase, stage, v1, mt * TM + q * 32, col0, lane);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[as]);
      if (++as == 2) { as = 0; aph ^= 1; }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // every bulk store of this warp has landed
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

// ------------------------------------------------------------------ step 2b: 256 x 256 tile variant
// Why: the 128x256 kernel needs 96 KiB of operands per 1536 tensor-cycles = 62.5 B/clk per SM, which is the
// SM's whole L2 read port; ncu showed the tensor pipe idle 26 % of a tile waiting for TMA
// (profiles/r01_gemm_ncu.md).  A 256x256 tile (two M=128 MMAs sharing each B slice) needs
// 64 KiB per 1536 cycles = 42.7 B/clk.  k-block = 16 floats (64-byte rows, SWIZZLE_64B) so that
// 3 stages x 64 KiB fit; the two 128x256 accumulators fill TMEM (512 columns), so the epilogue is
// not overlapped with the next tile's MMAs (<= 2 % at K = 4096).
//
// NB = 192 is the same kernel with 256 x 192 tiles (48.6 B/clk): for N = 768 (GPT-2's projections) 8192 rows make
// 128 tiles instead of 96 on 148 SMs.
constexpr int T2M = 256, T2K = 16;
constexpr int T2_PLANE = 256 * T2K * 4;                 // A plane, 16 KiB: 256 rows x 64 B
__host__ __device__ constexpr int t2_bplane(int nb) { return nb * T2K * 4; }
__host__ __device__ constexpr int t2_stage_bytes(int nb) { return 2 * T2_PLANE + 2 * t2_bplane(nb); }   // A_hi A_lo B_hi B_lo
__host__ __device__ constexpr int t2_smem_bytes(int nb, int nst) { return nst * t2_stage_bytes(nb) + 1024 + EPI_BYTES + 256; }
static_assert(t2_smem_bytes(256, 3) <= 232448 && t2_stage_bytes(192) % 1024 == 0 && t2_stage_bytes(256) % 1024 == 0, "shared-memory budget");
__host__ __device__ constexpr uint32_t idesc_tf32(int n) { return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24); }

template <int NB, int NST>
__global__ void __launch_bounds__(EPI8_THREADS, 1)
gemm_tc256_kernel(const __grid_constant__ CUtensorMap map_ah, const __grid_constant__ CUtensorMap map_al,
                  const __grid_constant__ CUtensorMap map_bh, const __grid_constant__ CUtensorMap map_bl,
                  const __grid_constant__ CUtensorMap map_c, const __grid_constant__ TcArgs a) {
  constexpr int T2N = NB, T2_STAGES = NST, B_PLANE = t2_bplane(NB), T2_STAGE_BYTES = t2_stage_bytes(NB);
  constexpr uint32_t kIdesc = idesc_tf32(NB), kIdescB = idesc_bf16_mn(128, NB);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-byte aligned, and still known to be SHARED (LDS / STS, not generic LD / ST)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + T2_STAGES * T2_STAGE_BYTES + EPI_BYTES);
  uint64_t* full = bars;                       // [T2_STAGES]
  uint64_t* empty = bars + T2_STAGES;          // [T2_STAGES]
  uint64_t* acc_full = bars + 2 * T2_STAGES;   // [1]
  uint64_t* acc_empty = acc_full + 1;          // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = a.tiles_m * a.tiles_n * a.batch;

  if (threadIdx.x == 0) {
    for (int s = 0; s < T2_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, EPI8_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      const uint32_t tx_bytes = a.passes != 1 ? T2_STAGE_BYTES : T2_PLANE + B_PLANE;
      int s = 0; uint32_t ph = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        int b, mt, nt;
        tile_coords(t, a, b, mt, nt);
        const int ba = a.a_bcast ? 0 : b, bb = a.b_bcast ? 0 : b;
        for (int kb = 0; kb < a.nkb; ++kb) {
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* st = smem + s * T2_STAGE_BYTES;
          mbar_expect_tx(&full[s], tx_bytes);
          tma_load_3d(&map_ah, &full[s], st, kb * T2K, mt * T2M, ba);
          tma_load_3d(&map_bh, &full[s], st + 2 * T2_PLANE, kb * T2K, nt * T2N, bb);
          if (a.passes != 1) {
            tma_load_3d(&map_al, &full[s], st + T2_PLANE, kb * T2K, mt * T2M, ba);
            tma_load_3d(&map_bl, &full[s], st + 2 * T2_PLANE + B_PLANE, kb * T2K, nt * T2N, bb);
          }
          if (++s == T2_STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      int s = 0; uint32_t ph = 0;
      uint32_t aph = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        mbar_wait(acc_empty, aph ^ 1);          // both accumulators drained by the epilogue
        tc_fence_after();
        for (int kb = 0; kb < a.nkb; ++kb) {
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t st = smem_u32(smem + s * T2_STAGE_BYTES);
          const uint64_t dbh = make_desc_sw64(st + 2 * T2_PLANE), dbl = make_desc_sw64(st + 2 * T2_PLANE + B_PLANE);
#pragma unroll
          for (int h = 0; h < 2; ++h) {         // rows [0,128) and [128,256) of the A tile: + 128 rows x 64 B
            const uint64_t dah = make_desc_sw64(st + h * (128 * 64)), dal = make_desc_sw64(st + T2_PLANE + h * (128 * 64));
            const uint32_t d = tmem_base + (uint32_t)(h * 256);
            if (a.passes == 2) {
              // the 64-byte swizzle row of the mixed plane: hb[0:16] then lb[0:16]; one bf16 MMA (K = 16) per correction
              tc_mma_bf16(d, dal, dbl + 2, kIdescB, kb != 0);
              tc_mma_bf16(d, dal + 2, dbl, kIdescB, 1);
            }
#pragma unroll
            for (int k = 0; k < T2K / UK; ++k) {
              const uint64_t adv = (uint64_t)((k * UK * 4) >> 4);
              if (a.passes == 2) {
                tc_mma_tf32(d, dah + adv, dbh + adv, kIdesc, 1);
              } else if (a.passes == 3) {
                tc_mma_tf32(d, dal + adv, dbh + adv, kIdesc, (kb | k) != 0);
                tc_mma_tf32(d, dah + adv, dbl + adv, kIdesc, 1);
                tc_mma_tf32(d, dah + adv, dbh + adv, kIdesc, 1);
              } else {
                tc_mma_tf32(d, dah + adv, dbh + adv, kIdesc, (kb | k) != 0);
              }
            }
          }
          tc_commit(&empty[s]);
          if (++s == T2_STAGES) { s = 0; ph ^= 1; }
        }
        tc_commit(acc_full);
        aph ^= 1;
      }
    }
  } else {
    const int q = warp & 3, cgrp = (warp - 2) >> 2;       // TMEM lane quarter (fixed by warp % 4), column half of the tile
    constexpr int NCH = T2N / 64;                         // 32-column chunks per warp: 4 (256-column tiles) or 3 (192)
    uint8_t* blk = smem + T2_STAGES * T2_STAGE_BYTES + (warp - 2) * EPI8_SLOT;
    uint32_t aph = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      int b, mt, nt;
      tile_coords(t, a, b, mt, nt);
      mbar_wait(acc_full, aph);
      tc_fence_after();
      float* cbase = a.C + (int64_t)b * a.sc;
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * 256 + cgrp * (T2N / 2));
        const int row0 = mt * T2M + h * 128 + q * 32;
        uint32_t v0[32], v1[32];
        tc_ld32(tbase, v0);
#pragma unroll 1
        for (int c = 0; c < NCH; c += 2) {
          tc_ld_wait();
          if (c + 1 < NCH) tc_ld32(tbase + (c + 1) * 32, v1);
          int col0 = nt * T2N + cgrp * (T2N / 2) + c * 32;
          if (col0 < a.n) tc_epilogue32(a, v0, row0, col0, lane);
          if (a.c_tma) { tc_store32_tma1(a, &map_c, blk, v0, row0, col0, b, lane); tc_plane32(a, blk, v0, row0, col0, b, lane); }
          else tc_store32_rows(a, cbase, v0, row0, col0, lane);
          if (c + 1 < NCH) {
            tc_ld_wait();
            if (c + 2 < NCH) tc_ld32(tbase + (c + 2) * 32, v0);
            col0 += 32;
            if (col0 < a.n) tc_epilogue32(a, v1, row0, col0, lane);
            if (a.c_tma) { tc_store32_tma1(a, &map_c, blk, v1, row0, col0, b, lane); tc_plane32(a, blk, v1, row0, col0, b, lane); }
            else tc_store32_rows(a, cbase, v1, row0, col0, lane);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty);
      aph ^= 1;
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

}  // namespace wf2
#include "gemm_tc2sm.cuh"   // step 2b': the same plane-fed loop on a CTA pair (cta_group::2), half of B per CTA
#include "gemm_raw.cuh"   // step 2c: the same main loop on RAW fp32 tiles (no operand pre-pass) + the attention-score epilogues
namespace wf2 {

// ------------------------------------------------------------------ step 3: fused small-problem kernel
// The MLP training step's GEMMs (C4: 1024x256x784 ... 10x256x1024) are 0.005-0.4 GFLOP each: too small for
// two extra launches that write hi/lo planes, and too few 128x256 tiles to fill 148 SMs.  Here ONE launch
// does everything for a 128 x 64 tile and one K slice:
//   all 8 warps : load the fp32 operands straight from global memory (any transpose flags, bounds and
//                 alignment), split them into TF32 hi/lo in registers and store the four planes into
//                 shared memory in the K-major SWIZZLE_128B layout the UMMA descriptors expect
//                 (generic-proxy stores + fence.proxy.async), double buffered, next block's global
//                 loads in flight under the current MMAs;
//   one thread  : 3 x tcgen05.mma kind::tf32 128x64x8 per 8-wide k step into a 64-column TMEM accumulator;
//   warps 0-3   : tcgen05.ld the accumulator into a shared-memory tile;
//   split-K     : the K axis is spread over a thread-block cluster (<= 8 CTAs); partial tiles are folded
//                 through distributed shared memory in slice order, and the fused epilogue (bias / relu /
//                 relu-mask) is applied on the way out.
constexpr int FM = 128, FN = 64;
constexpr int F_A_PLANE = FM * TK * 4;                    // 16 KiB
constexpr int F_B_PLANE = FN * TK * 4;                    // 8 KiB
constexpr int F_STAGE = 2 * F_A_PLANE + 2 * F_B_PLANE;    // 48 KiB
constexpr int F_STAGES = 2;
constexpr int F_PLD = FN + 4;                             // padded row of the partial tile (floats)
constexpr int F_SMEM = F_STAGES * F_STAGE + 1024 + 128;
constexpr uint32_t kIdescTf32F = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(FN >> 3) << 17) | ((uint32_t)(FM >> 4) << 24);
static_assert(FM * F_PLD * 4 <= F_STAGE, "partial tile must fit stage 0");

struct FusedGemmArgs {
  const float* A; const float* B; float* C;
  int64_t lda, ldb, ldc;
  int64_t sa, sb, sc;        // batch strides (elements); blockIdx.z = batch index * ksplit + K slice
  int m, n, k;
  int ta, tb;
  int a_vec, b_vec, c_vec;   // operand / result rows are 16-byte aligned: 128-bit global accesses
  int ksplit, kchunk;        // cluster size along z, K per slice (multiple of 32)
  const float* bias; const float* mask; int64_t ldmask; int act;
  int a_hint;                // cache hint for the streamed K-major A rows (0 none, 1 L2::256B, 2 + L1::no_allocate, 3 L1::no_allocate)
};

// byte offset of element (r, k) inside a K-major SWIZZLE_128B plane (128-byte rows, 16-byte chunks XOR r%8)
__device__ __forceinline__ uint32_t sw128_off(int r, int k) {
  return (uint32_t)(r * 128 + ((((k >> 2) ^ (r & 7)) << 4) | ((k & 3) << 2)));
}

// Per-thread view of one operand tile (ROWS x 32 per k-block), fixed for the whole K slice: thread t owns
// ROWS/32 16-byte chunks.  kmajor (element (r, k) at src[r*ld + k]): a chunk is 4 consecutive k of one row;
// otherwise ((r, k) at src[k*ld + r]) it is 4 consecutive rows of one k, lanes running along k so that the
// transposing shared-memory stores are conflict free.  Pointers, swizzled shared-memory offsets and
// validity are computed once; a k-block then costs one add per chunk.
template <int ROWS>
struct FTile {
  const float* p[ROWS / 32];     // this chunk's first element in the slice's first k-block
  int nv[ROWS / 32];             // valid elements along the chunk's row axis: kmajor 4 or 0, else 0..4
  // (k of the chunk's first element inside a k-block and its shared-memory offset are recomputed from
  //  threadIdx: a few bit operations, cheaper than 12 more live registers)
  __device__ __forceinline__ int kk_of(int it) const { const int q = threadIdx.x + it * 256; return kmajor ? 4 * (q & 7) : (q & 31); }
  int64_t step;                  // elements to advance per k-block
  bool kmajor, vec;
  __device__ __forceinline__ void init(const float* src, int64_t ld, bool kmaj, bool v, int r0, int rmax, int kbeg) {
    kmajor = kmaj; vec = v;
    step = kmaj ? TK : (int64_t)TK * ld;
#pragma unroll
    for (int it = 0; it < ROWS / 32; ++it) {
      const int q = threadIdx.x + it * 256;
      if (kmaj) {
        const int c = q & 7, r = q >> 3, gr = r0 + r;
        nv[it] = gr < rmax ? 4 : 0;
        p[it] = src + (int64_t)(gr < rmax ? gr : 0) * ld + kbeg + 4 * c;
      } else {
        const int k = q & 31, rq = q >> 5, gr = r0 + 4 * rq;
        nv[it] = gr >= rmax ? 0 : (rmax - gr >= 4 ? 4 : rmax - gr);
        p[it] = src + (int64_t)(kbeg + k) * ld + (gr < rmax ? gr : 0);
      }
    }
  }
  // blk: k-block index inside the slice; krem = kend - (kbeg + blk*32) (> 0)
  __device__ __forceinline__ void load(float4 (&reg)[ROWS / 32], int blk, int krem) const {
    const bool fullk = krem >= TK;
#pragma unroll
    for (int it = 0; it < ROWS / 32; ++it) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      const float* q = p[it] + blk * step;
      const int kk = kk_of(it);
      if (nv[it] == 4 && vec && (fullk || (kmajor ? kk + 3 < krem : kk < krem))) {
        v = *reinterpret_cast<const float4*>(q);
      } else if (nv[it] > 0) {
        if (kmajor) {
          const int left = krem - kk;
          if (left > 0) v.x = q[0];
          if (left > 1) v.y = q[1];
          if (left > 2) v.z = q[2];
          if (left > 3) v.w = q[3];
        } else if (kk < krem) {
          v.x = q[0];
          if (nv[it] > 1) v.y = q[1];
          if (nv[it] > 2) v.z = q[2];
          if (nv[it] > 3) v.w = q[3];
        }
      }
      reg[it] = v;
    }
  }
  // interior tile, aligned operands, whole k-blocks (uniform over the CTA): one 128-bit load per chunk, no tests
  __device__ __forceinline__ void load_fast(float4 (&reg)[ROWS / 32], int blk, int hint = 0) const {
#pragma unroll
    for (int it = 0; it < ROWS / 32; ++it) {
      const float* q = p[it] + blk * step;
      if (hint == 1)
        asm volatile("ld.global.L2::256B.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(reg[it].x), "=f"(reg[it].y), "=f"(reg[it].z), "=f"(reg[it].w) : "l"(q));
      else if (hint == 2)
        asm volatile("ld.global.L1::no_allocate.L2::256B.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(reg[it].x), "=f"(reg[it].y), "=f"(reg[it].z), "=f"(reg[it].w) : "l"(q));
      else if (hint == 3)
        asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(reg[it].x), "=f"(reg[it].y), "=f"(reg[it].z), "=f"(reg[it].w) : "l"(q));
      else
        reg[it] = *reinterpret_cast<const float4*>(q);
    }
  }
  // hi = x with the 13 low mantissa bits cleared (a TF32 number), lo = x - hi (exact).  lo is stored as is: the
  // tensor core reads 19 bits of it, an error of 2^-10 |lo| <= 2^-20 |x|, the same order as the split itself.
  __device__ __forceinline__ static void split(float x, float& hi, float& lo) {
    hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
    lo = (fabsf(x) <= 3.0e38f) ? x - hi : 0.f;           // inf / nan live in hi only
  }
  __device__ __forceinline__ void store(const float4 (&reg)[ROWS / 32], uint8_t* hi, uint8_t* lo) const {
#pragma unroll
    for (int it = 0; it < ROWS / 32; ++it) {
      float4 h, l;
      split(reg[it].x, h.x, l.x); split(reg[it].y, h.y, l.y); split(reg[it].z, h.z, l.z); split(reg[it].w, h.w, l.w);
      const int q = threadIdx.x + it * 256;
      if (kmajor) {
        const uint32_t o = sw128_off(q >> 3, 4 * (q & 7));
        *reinterpret_cast<float4*>(hi + o) = h;
        *reinterpret_cast<float4*>(lo + o) = l;
      } else {
        const float hh[4] = {h.x, h.y, h.z, h.w}, ll[4] = {l.x, l.y, l.z, l.w};
        const int k = q & 31, rq = q >> 5;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t o = sw128_off(4 * rq + j, k);
          *reinterpret_cast<float*>(hi + o) = hh[j];
          *reinterpret_cast<float*>(lo + o) = ll[j];
        }
      }
    }
  }
};

__device__ __forceinline__ float f_epilogue(const FusedGemmArgs& g, float v, int64_t gm, int64_t gn) {
  if (g.bias) v = __fadd_rn(v, g.bias[gn]);
  if (g.act == 1) v = __fmul_rn(v, v > 0.f ? 1.f : 0.f);
  if (g.mask) v = __fmul_rn(v, g.mask[gm * g.ldmask + gn] > 0.f ? 1.f : 0.f);
  return v;
}

// Main loop of the fused kernel.  A ring of GB = 4 k-blocks lives in registers: as soon as a block has been converted
// and stored, the load of the block GB steps ahead is issued into the same registers, so three blocks of global
// latency are always in flight (with the loads of a whole group issued together and then drained, GPT-2's P.V —
// K = 1024 per CTA — spent its time in long-scoreboard stalls: 260 us for 402 MB, profiles/r01d_attention_ncu.md).
template <bool FAST>
__device__ __forceinline__ void fused_mainloop(const FTile<FM>& ta_, const FTile<FN>& tb_, uint8_t* smem, uint64_t* freed,
                                               uint32_t tmem_base, int nkb, int kbeg, int kend, uint32_t& ph0, uint32_t& ph1, int hint) {
  constexpr int GB = 4;
  float4 ra[GB][FM / 32], rb[GB][FN / 32];
#pragma unroll
  for (int u = 0; u < GB; ++u)
    if (u < nkb) {
      if (FAST) { ta_.load_fast(ra[u], u, hint); tb_.load_fast(rb[u], u); }
      else { const int krem = kend - (kbeg + u * TK); ta_.load(ra[u], u, krem); tb_.load(rb[u], u, krem); }
    }
#pragma unroll 1
  for (int kg = 0; kg < nkb; kg += GB) {
#pragma unroll
    for (int u = 0; u < GB; ++u) {
      const int kb = kg + u;
      if (kb < nkb) {                                   // uniform across the CTA
        const int s = kb & 1;
        uint8_t* st = smem + s * F_STAGE;
        if (kb >= F_STAGES) {                           // the MMAs of block kb-2 must be done reading this stage
          if (s == 0) { mbar_wait(&freed[0], ph0); ph0 ^= 1; } else { mbar_wait(&freed[1], ph1); ph1 ^= 1; }
          tc_fence_after();
        }
        ta_.store(ra[u], st, st + F_A_PLANE);
        tb_.store(rb[u], st + 2 * F_A_PLANE, st + 2 * F_A_PLANE + F_B_PLANE);
        if (kb + GB < nkb) {                            // refill this ring slot
          if (FAST) { ta_.load_fast(ra[u], kb + GB, hint); tb_.load_fast(rb[u], kb + GB); }
          else { const int krem = kend - (kbeg + (kb + GB) * TK); ta_.load(ra[u], kb + GB, krem); tb_.load(rb[u], kb + GB, krem); }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the tensor core
        tc_fence_before();
        __syncthreads();
        if (threadIdx.x == 0) {   // (elect_one() here removes the per-MMA ELECT loops but measured SLOWER: C4 step 90.4 -> 92.3 us)
          tc_fence_after();
          const uint32_t sa = smem_u32(st);
          const uint64_t dah = make_desc_sw128(sa), dal = make_desc_sw128(sa + F_A_PLANE);
          const uint64_t dbh = make_desc_sw128(sa + 2 * F_A_PLANE), dbl = make_desc_sw128(sa + 2 * F_A_PLANE + F_B_PLANE);
#pragma unroll
          for (int k = 0; k < TK / UK; ++k) {
            const uint64_t adv = (uint64_t)((k * UK * 4) >> 4);
            tc_mma_tf32(tmem_base, dal + adv, dbh + adv, kIdescTf32F, (kb | k) != 0);
            tc_mma_tf32(tmem_base, dah + adv, dbl + adv, kIdescTf32F, 1);
            tc_mma_tf32(tmem_base, dah + adv, dbh + adv, kIdescTf32F, 1);
          }
          tc_commit(&freed[s]);
        }
      }
    }
  }
}

__global__ void __launch_bounds__(256, 2) gemm_tc_fused_kernel(const __grid_constant__ FusedGemmArgs g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-byte aligned, and still known to be SHARED (LDS / STS, not generic LD / ST)
  uint64_t* freed = reinterpret_cast<uint64_t*>(smem + F_STAGES * F_STAGE);   // [2] MMAs that read stage s have retired
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(freed + F_STAGES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ks = (int)(blockIdx.z % g.ksplit);          // cluster = ksplit consecutive z: rank in cluster == K slice
  const int64_t bz = blockIdx.z / g.ksplit;             // batch index
  const int m0 = blockIdx.y * FM, n0 = blockIdx.x * FN;
  const int kbeg = ks * g.kchunk;
  const int kend = min(g.k, kbeg + g.kchunk);
  const int nkb = kend > kbeg ? (kend - kbeg + TK - 1) / TK : 0;
  const bool a_kmajor = !g.ta, b_kmajor = g.tb != 0;    // A is (m k) row-major unless transposed; B^T is (n k) when tb

  if (threadIdx.x == 0) {
    for (int s = 0; s < F_STAGES; ++s) mbar_init(&freed[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(64) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // interior tiles of aligned operands (every tile of GPT-2's P.V, most of the MLP's) skip the per-chunk tests
  FTile<FM> ta_;
  FTile<FN> tb_;
  ta_.init(g.A + bz * g.sa, g.lda, a_kmajor, g.a_vec != 0, m0, g.m, kbeg);
  tb_.init(g.B + bz * g.sb, g.ldb, b_kmajor, g.b_vec != 0, n0, g.n, kbeg);
  uint32_t ph0 = 0, ph1 = 0;                            // parity of freed[0] / freed[1]
  const bool fast = g.a_vec && g.b_vec && m0 + FM <= g.m && n0 + FN <= g.n && nkb > 0 && ((kend - kbeg) % TK) == 0;
  if (fast) fused_mainloop<true>(ta_, tb_, smem, freed, tmem_base, nkb, kbeg, kend, ph0, ph1, g.a_hint);
  else      fused_mainloop<false>(ta_, tb_, smem, freed, tmem_base, nkb, kbeg, kend, ph0, ph1, 0);
  // ---- all MMAs retired?  (the last commit covers every earlier one)
  if (nkb > 0) {
    // commits retire in order, so waiting for the last block's commit covers every MMA; the other
    // stage's last phase is consumed too so that no arrival is left pending
    if (nkb >= 2) { mbar_wait(&freed[0], ph0); mbar_wait(&freed[1], ph1); }
    else mbar_wait(&freed[0], ph0);
    tc_fence_after();
  }
  __syncthreads();
  // ---- accumulator -> shared-memory partial tile Ps[128][F_PLD] (stage 0 is free now)
  float* Ps = reinterpret_cast<float*>(smem);
  if (warp < 4) {
    const int row = warp * 32 + lane;
    float* pr = Ps + row * F_PLD;
    if (nkb > 0) {
      const uint32_t tbase = tmem_base + ((uint32_t)(warp * 32) << 16);
#pragma unroll
      for (int c = 0; c < FN / 32; ++c) {
        uint32_t v[32];
        tc_ld32(tbase + c * 32, v);
        tc_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<float4*>(pr + c * 32 + 4 * j) =
              make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
      }
    } else {
#pragma unroll
      for (int j = 0; j < FN / 4; ++j) *reinterpret_cast<float4*>(pr + 4 * j) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  tc_fence_before();
  // ---- fold the K slices (cluster) and write C
  const int ksplit = g.ksplit;
  if (ksplit > 1) {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  } else {
    __syncthreads();
  }
  const int rows_per = FM / ksplit;                     // this CTA writes rows [ks*rows_per, (ks+1)*rows_per) of the tile
  for (int e = threadIdx.x; e < rows_per * (FN / 4); e += 256) {
    const int r = ks * rows_per + e / (FN / 4), c4 = (e % (FN / 4)) * 4;
    const uint32_t local = smem_u32(Ps + r * F_PLD + c4);
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int p = 0; p < ksplit; ++p) {
      float4 o;
      if (ksplit > 1) {
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(p));
        asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w) : "r"(remote) : "memory");
      } else {
        o = *reinterpret_cast<const float4*>(Ps + r * F_PLD + c4);
      }
      if (p == 0) t = o;
      else { t.x = __fadd_rn(t.x, o.x); t.y = __fadd_rn(t.y, o.y); t.z = __fadd_rn(t.z, o.z); t.w = __fadd_rn(t.w, o.w); }
    }
    const int64_t gm = m0 + r, gn = n0 + c4;
    if (gm < g.m) {
      float* cp = g.C + bz * g.sc + gm * g.ldc + gn;
      const float v[4] = {t.x, t.y, t.z, t.w};
      if (g.c_vec && gn + 3 < g.n && !g.bias && !g.act && !g.mask) {
        *reinterpret_cast<float4*>(cp) = t;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (gn + j < g.n) cp[j] = f_epilogue(g, v[j], gm, gn + j);
      }
    }
  }
  if (ksplit > 1)   // nobody leaves (or frees its shared memory) while a peer may still read its tile
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(64) : "memory");
  }
}

// ------------------------------------------------------------------ host side
static EncodeTiledFn g_encode = nullptr;

EncodeTiledFn tensor_map_encoder() {
  if (!g_encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
      set_error(WF2CU_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
      return nullptr;
    }
    g_encode = (EncodeTiledFn)fn;
  }
  return g_encode;
}
static float* g_ws = nullptr;            // hi/lo planes workspace (grow-only)
static size_t g_ws_bytes = 0;
static std::vector<void*> g_retired;     // older workspaces a captured graph may still reference
static bool g_attr_set = false;

void gemm_tc_const_release();
void planes_emit_release();
void gemm_tc_release() {
  planes_emit_release();
  if (g_ws) cudaFree(g_ws);
  for (void* p : g_retired) cudaFree(p);
  g_retired.clear();
  g_ws = nullptr; g_ws_bytes = 0; g_attr_set = false;
  gemm_tc_const_release();
}

static int ensure_ws(size_t bytes) {
  if (g_ws_bytes >= bytes) return 0;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(S().stream, &cs);
  if (cs != cudaStreamCaptureStatusNone)
    return -set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_gemm: the tensor-core workspace must grow (%zu bytes) while a CUDA graph is "
                      "being captured; run the same shapes once before wf2cu_graph_begin", bytes);
  if (g_ws) g_retired.push_back(g_ws);
  g_ws = nullptr; g_ws_bytes = 0;
  const size_t want = bytes + (bytes >> 2);
  cudaError_t e = cudaMalloc(&g_ws, want);
  if (e != cudaSuccess) return -cuda_fail(e, "cudaMalloc(gemm workspace)", __FILE__, __LINE__);
  g_ws_bytes = want;
  return 0;
}

// the planes workspace, for the other tcgen05 kernels of the library (attention.cu): nullptr + error set on failure
float* tc_workspace(size_t bytes) {
  if (ensure_ws(bytes)) return nullptr;
  return g_ws;
}

// ---- operand planes written by producers (wf2cu_emit_operand_planes) ---------------------------------------------
// The activations of a layer chain are the K-major A operand of the next MatmulNode: its pre-pass (read 4 + write 4 bytes per
// element, one launch) only restates what the producing kernel had in registers.  A producer that is asked to writes the mixed
// bf16 plane itself (GPT-2 small: LayerNorm -> QKV / fc, P.V -> proj, gelu(fc) -> proj: 4 pre-passes per block, 8 % of the
// forward).  One grow-only buffer, stream ordered like the workspace; retired buffers stay alive for captured graphs.
static float* g_emit2[2] = {nullptr, nullptr};   // alternate: the plane being consumed is never the one being written
static int g_emit_idx = 0;
static size_t g_emit_bytes = 0;
int gemm_tc_correction();

float* planes_emit_begin(int64_t rows, int64_t k) {
  S().emit_next = 0;
  if (rows < 1 || k < 32 || (k & 31) || S().branch != 0 || gemm_tc_correction() != 0) return nullptr;
  const size_t need = (size_t)rows * (size_t)k * 4;
  if (need > g_emit_bytes) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(S().stream, &cs);
    if (cs != cudaStreamCaptureStatusNone) return nullptr;      // no allocation inside a capture: the consumer runs its pre-pass
    // two buffers of the same size: a MatmulNode that consumes the plane its producer wrote may itself be asked to write one
    float* p[2] = {nullptr, nullptr};
    if (cudaMalloc(&p[0], need) != cudaSuccess || cudaMalloc(&p[1], need) != cudaSuccess) {
      cudaGetLastError();
      if (p[0]) cudaFree(p[0]);
      return nullptr;
    }
    for (int i = 0; i < 2; ++i) { if (g_emit2[i]) g_retired.push_back(g_emit2[i]); g_emit2[i] = p[i]; }
    g_emit_bytes = need;
  }
  g_emit_idx ^= 1;
  return g_emit2[g_emit_idx];
}

void planes_emit_release() {
  for (int i = 0; i < 2; ++i) { if (g_emit2[i]) cudaFree(g_emit2[i]); g_emit2[i] = nullptr; }
  g_emit_bytes = 0;
  S().fresh.ptr = nullptr; S().emit_next = 0;
}

void planes_emit_publish(const float* result, int64_t rows, int64_t k, int64_t ld, float* lo) {
  State::FreshPlanes& f = S().fresh;
  f.ptr = result; f.rows = rows; f.k = k; f.ld = ld; f.lo = lo;
}

static bool g_t256_auto = true;       // auto mode may take the 256x256-tile kernel (rule in gemm_tc_try)
// how the planes kernels form the correction terms hi.lo + lo.hi: 0 = two bf16 MMAs on mixed planes (default), 1 = two TF32
// MMAs on fp32 lo planes (classic 3xTF32).  WF2CU_GEMM_CORRECTION=tf32 selects 1 at start-up.
static int g_correction = -1;
int gemm_tc_correction() {
  if (g_correction < 0) { const char* e = getenv("WF2CU_GEMM_CORRECTION"); g_correction = (e && (e[0] == 't' || e[0] == '1')) ? 1 : 0; }
  return g_correction;
}
void gemm_tc_set_correction(int kind) { g_correction = kind ? 1 : 0; }

static int make_map256(CUtensorMap* map, float* base, int rows, int Kp, int nbatch, int box_rows) {
  cuuint64_t dims[3] = {(cuuint64_t)Kp, (cuuint64_t)rows, (cuuint64_t)nbatch};
  cuuint64_t strides[2] = {(cuuint64_t)Kp * 4, (cuuint64_t)rows * Kp * 4};
  cuuint32_t box[3] = {(cuuint32_t)T2K, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return -set_error(WF2CU_ERR_CUDA, "cuTensorMapEncodeTiled(256x16, SWIZZLE_64B) failed (%d) rows=%d Kp=%d batch=%d", (int)r, rows, Kp, nbatch);
  return 0;
}

// C as (n, m, batch) for the epilogue's bulk tensor stores: 32 x 32 boxes, SWIZZLE_128B.  Returns 1 when the map was
// built, 0 when C's layout does not fit a tensor map (the epilogue then stores from registers).
static int g_c_tma = -1;              // WF2CU_GEMM_CTMA=0 turns the bulk-store epilogue off
static int make_map_c(CUtensorMap* map, float* C, int64_t m, int64_t n, int64_t ldc, int64_t sc, int64_t batch) {
  memset(map, 0, sizeof(*map));
  if (g_c_tma < 0) { const char* e = getenv("WF2CU_GEMM_CTMA"); g_c_tma = (e && e[0] == '0') ? 0 : 1; }
  if (!g_c_tma) return 0;
  if ((reinterpret_cast<uintptr_t>(C) & 15) != 0 || (ldc & 3) != 0 || ldc < n) return 0;
  int64_t bstride = sc;
  if (batch == 1) bstride = m * ldc;
  else if (sc < m * ldc || (sc % ldc) != 0) return 0;       // overlapping or oddly pitched batches: keep the plain path
  cuuint64_t dims[3] = {(cuuint64_t)n, (cuuint64_t)m, (cuuint64_t)batch};
  cuuint64_t strides[2] = {(cuuint64_t)ldc * 4, (cuuint64_t)bstride * 4};
  cuuint32_t box[3] = {32, 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, C, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 1 : 0;
}

static int launch_split(const float* src, float* hi, float* lo, int64_t ld, int64_t sx, int R, int K, int Kp, int nbatch,
                        bool r_contiguous, int bf_blk, bool raw_hi = false) {
  SplitArgs sa{src, hi, lo, ld, sx, R, K, Kp, 0, bf_blk, raw_hi ? 1 : 0};
  if (!r_contiguous) {
    sa.vec = ((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (ld & 3) == 0 && (sx & 3) == 0) ? 1 : 0;
    const int64_t work = (int64_t)R * (Kp >> 2);
    dim3 grid((unsigned)((work + 255) / 256), (unsigned)nbatch);
    if (raw_hi) split_planes_direct_kernel<true><<<grid, 256, 0, S().stream>>>(sa);
    else        split_planes_direct_kernel<false><<<grid, 256, 0, S().stream>>>(sa);
  } else {
    dim3 grid((unsigned)((R + 31) / 32), (unsigned)(Kp / 32), (unsigned)nbatch);
    split_planes_transpose_kernel<<<grid, 256, 0, S().stream>>>(sa);
  }
  const int rc = post_launch("wf2cu_gemm(split)");
  return rc ? -rc : 0;
}

// A K-major operand plane or the operand ITSELF as (k, rows, batch) with a row pitch of `ld` floats; k beyond the true extent is
// zero-filled by the TMA unit (OOB fill), so a raw source needs no padded copy.
static int make_map_k(CUtensorMap* map, const float* base, int64_t kdim, int64_t rows, int64_t ld, int64_t nbatch, int64_t bstride,
                      int box_k, int box_rows) {
  cuuint64_t dims[3] = {(cuuint64_t)kdim, (cuuint64_t)rows, (cuuint64_t)(nbatch < 1 ? 1 : nbatch)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 4, (cuuint64_t)(nbatch > 1 ? bstride : rows * ld) * 4};
  cuuint32_t box[3] = {(cuuint32_t)box_k, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        box_k == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return -set_error(WF2CU_ERR_CUDA, "cuTensorMapEncodeTiled(operand, box %dx%d) failed (%d) k=%lld rows=%lld ld=%lld batch=%lld",
                                           box_k, box_rows, (int)r, (long long)kdim, (long long)rows, (long long)ld, (long long)nbatch);
  return 0;
}

static bool raw_operand_ok(const float* p, int64_t ld, int64_t bstride, int64_t nbatch);

// WF2CU_GEMM_RAWHI=0: always write RN hi planes (A/B measurements; the round-1 behaviour)
static bool rawhi_enabled() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("WF2CU_GEMM_RAWHI"); on = (e && e[0] == '0') ? 0 : 1; }
  return on != 0;
}

// ---- prepared (constant) operands ------------------------------------------------------------------------------
// BLAS libraries offer packed-operand entry points (sgemm_pack / sgemm_compute) for a matrix that is multiplied many times; here
// the host declares a storage range constant (wf2cu_gemm_const_operand: the parameters of a model in inference) and the planes
// of an operand that starts inside such a range are split ONCE into buffers of their own and found again by
// (pointer, pitch, shape, layout).  An upload / memset / free that touches the range marks the planes stale (re-split on the
// next product) or drops them.  GPT-2 small: 48 weight matrices, 680 MB of planes, 48 transposing split launches per forward saved.
struct ConstRange { uintptr_t lo, hi; };
struct PlaneKey {
  const float* p; int64_t ld, sx; int R, K, Kp, nb, rcontig, bf_blk, raw_hi;
  bool operator<(const PlaneKey& o) const {
    return std::tie(p, ld, sx, R, K, Kp, nb, rcontig, bf_blk, raw_hi) < std::tie(o.p, o.ld, o.sx, o.R, o.K, o.Kp, o.nb, o.rcontig, o.bf_blk, o.raw_hi);
  }
};
struct PlaneVal { float* hi; float* lo; size_t bytes; uintptr_t src_lo, src_hi; bool stale; };
static std::vector<ConstRange> g_const;
static std::map<PlaneKey, PlaneVal> g_planes;
static size_t g_planes_bytes = 0;

static bool in_const_range(const float* p) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  for (const ConstRange& r : g_const)
    if (a >= r.lo && a < r.hi) return true;
  return false;
}

static void drop_plane(PlaneVal& v) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(S().stream, &cs);
  if (cs != cudaStreamCaptureStatusNone) { if (v.hi) g_retired.push_back(v.hi); if (v.lo) g_retired.push_back(v.lo); }
  else { if (v.hi) cudaFree(v.hi); if (v.lo) cudaFree(v.lo); }
  g_planes_bytes -= v.bytes;
}

// the storage [ptr, ptr + bytes) was written (stale = true: keep the buffers, split again) or released (drop them)
void gemm_tc_const_touch(const void* ptr, size_t bytes, bool released) {
  {
    // a transfer into (or the release of) the tensor a producer wrote planes for: the planes no longer describe it
    const State::FreshPlanes& f = S().fresh;
    if (f.ptr) {
      const uintptr_t a0 = reinterpret_cast<uintptr_t>(ptr), a1 = a0 + bytes;
      const uintptr_t f0 = reinterpret_cast<uintptr_t>(f.ptr), f1 = f0 + (size_t)f.rows * (size_t)f.ld * 4;
      if (a0 < f1 && f0 < a1) S().fresh.ptr = nullptr;
    }
  }
  if (g_planes.empty() && (!released || g_const.empty())) return;
  const uintptr_t lo = reinterpret_cast<uintptr_t>(ptr), hi = lo + bytes;
  for (auto it = g_planes.begin(); it != g_planes.end();) {
    if (it->second.src_lo < hi && lo < it->second.src_hi) {
      if (released) { drop_plane(it->second); it = g_planes.erase(it); continue; }
      it->second.stale = true;
    }
    ++it;
  }
  if (released)
    for (size_t i = 0; i < g_const.size();)
      if (g_const[i].lo < hi && lo < g_const[i].hi) g_const.erase(g_const.begin() + i); else ++i;
}

int gemm_tc_const_declare(const void* ptr, size_t bytes, int on) {
  const uintptr_t lo = reinterpret_cast<uintptr_t>(ptr), hi = lo + bytes;
  if (on) {
    for (const ConstRange& r : g_const)
      if (r.lo <= lo && hi <= r.hi) return 0;
    g_const.push_back(ConstRange{lo, hi});
    return 0;
  }
  gemm_tc_const_touch(ptr, bytes, true);
  return 0;
}

void gemm_tc_const_release() {
  for (auto& kv : g_planes) { if (kv.second.hi) cudaFree(kv.second.hi); if (kv.second.lo) cudaFree(kv.second.lo); }
  g_planes.clear(); g_const.clear(); g_planes_bytes = 0;
}

void gemm_tc_const_stats(int64_t* entries, int64_t* bytes) { *entries = (int64_t)g_planes.size(); *bytes = (int64_t)g_planes_bytes; }

// the planes of a constant operand: found, or allocated here (fresh = they still have to be filled).  <0 = -(error code)
static int const_planes(const PlaneKey& key, size_t plane_bytes, size_t src_floats, PlaneVal** out, bool* fresh) {
  auto it = g_planes.find(key);
  if (it != g_planes.end()) { *out = &it->second; *fresh = it->second.stale; it->second.stale = false; return 0; }
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(S().stream, &cs);
  if (cs != cudaStreamCaptureStatusNone)
    return -set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_gemm: the planes of a constant operand must be allocated (%zu bytes) while a CUDA graph is "
                      "being captured; run the same product once before wf2cu_graph_begin", plane_bytes);
  PlaneVal v{nullptr, nullptr, 0, reinterpret_cast<uintptr_t>(key.p), reinterpret_cast<uintptr_t>(key.p) + src_floats * 4, false};
  cudaError_t e = cudaSuccess;
  if (!key.raw_hi) e = cudaMalloc(&v.hi, plane_bytes);
  if (e == cudaSuccess) e = cudaMalloc(&v.lo, plane_bytes);
  if (e != cudaSuccess) { if (v.hi) cudaFree(v.hi); return -cuda_fail(e, "cudaMalloc(constant operand planes)", __FILE__, __LINE__); }
  v.bytes = (key.raw_hi ? 1 : 2) * plane_bytes;
  g_planes_bytes += v.bytes;
  *out = &g_planes.emplace(key, v).first->second;
  *fresh = true;
  return 0;
}

// auto rule for the 256-row tile kernels (0 = stay on the 128x256 kernel): they cannot overlap the epilogue with the
// next tile, so they want a long K, and they have fewer tiles, so they want enough of them.
// Cost model fitted on B200 (scripts/gemm_t256.py, gpurun_out/gemm_t256.log), in units of one 128x256 MMA column of
// depth 1: waves x (MMA time of a tile / its efficiency + the exposed epilogue).  The 128x256 kernel is L2-port bound
// (x1.3) but hides its epilogue; 256x256 runs at the tensor rate; 256x192 pays x1.12 for its smaller B slices.
//   8192x768x3072: 286 / 242 / 212 us (128x256 / 256x256 / 256x192)   8192x768x768: 82 / 78 / 68
//   8192x1536x1536: 207 / 230 / 196     8192x2304x768: 139 / 136 / 156     2048^3: 103 / 154 / 131     4096^3: 679 / 577 / 680
static int pick_tile256(int64_t kp, int64_t tiles128, int64_t tiles256, int64_t tiles192, int64_t tiles_pair, int sms, bool bf16_corr) {
  auto waves = [&](int64_t t) { return (double)((t + sms - 1) / sms); };
  auto pair_waves = [&](int64_t t) { const int64_t pairs = sms / 2; return (double)((t + pairs - 1) / pairs); };
  double c128, c256, c192, cpair;
  if (bf16_corr) {
    // TF32 + bf16 corrections (r02d, scripts/gemm_hybrid_bench.py): 35 / 53 / 44.5 ns per k per tile; 256-row epilogues ~8 / 6 us
    //   4096^3: 624 / 490 / 606 us   8192^3: 3944 / 3217 / 3814   2048^3: 95 / 119 / 112   8192x2304x768: 124 / 112 / 134
    //   8192x768x768: 72 / 66 / 62   8192x3072x768: 171 / 154 / 170   8192x768x3072: 262 / 202 / 193
    // CTA-pair kernel (512 x 256 pair tiles on sms / 2 pairs, r02r, scripts/gemm_pair_bench.py): 4096^3 446 (NT 420), 8192^3 3125,
    //   8192x2304x768 107, 8192x768x768 64, 8192x3072x768 144, 8192x768x3072 179, 2048^3 117
    // refitted with eight epilogue warps in the 256-row kernels (r03R, scripts/gemm_gpt2_shapes.py with constant weights, kernel only):
    //   8192x2304x768: 113 / 100 / 120 / 94 us (128x256 / 256x256 / 256x192 / pair)   8192x768x768: 66 / 57 / 52 / 54
    //   8192x3072x768: 161 / 140 / 152 / 131   8192x768x3072: 252 / 189 / 177 / 164   4096^3: 603 / 463 / 570 / 428   2048^3: 88 / 112 / 103 / 111
    c128 = waves(tiles128) * (1.02 * (double)kp + 40.0);
    c256 = waves(tiles256) * (1.36 * (double)kp + 315.0);
    c192 = waves(tiles192) * (1.26 * (double)kp + 271.0);
    cpair = pair_waves(tiles_pair) * (1.09 * (double)kp + 449.0);
  } else {
    c128 = waves(tiles128) * 1.30 * (double)kp;
    c256 = waves(tiles256) * (2.00 * (double)kp + 450.0);
    c192 = waves(tiles192) * (1.68 * (double)kp + 337.0);
    cpair = pair_waves(tiles_pair) * (1.90 * (double)kp + 450.0);
  }
  if (tiles_pair <= 0) cpair = 1e300;
  if (c128 <= c256 * 1.02 && c128 <= c192 * 1.02 && c128 <= cpair * 1.02) return 0;
  if (cpair <= c256 && cpair <= c192) return 512;
  return c256 <= c192 ? 256 : 192;
}

// WF2CU_GEMM_PAIR=0: auto never takes the CTA-pair kernel; =1: always when the shape allows (A/B measurements)
static int pair_override() {
  static int force = -2;
  if (force == -2) { const char* e = getenv("WF2CU_GEMM_PAIR"); force = e ? atoi(e) : -1; }
  return force;
}

// returns 1 when it handled the problem, 0 when it declines, <0 = -(error code)
int gemm_tc_try(int ta, int tb, int64_t m, int64_t n, int64_t k, const float* A, int64_t lda, const float* B, int64_t ldb,
                float* C, int64_t ldc, int64_t batch, int64_t sa, int64_t sb, int64_t sc, int mode, const float* bias, int act,
                const float* res, int64_t ldres) {
  // mode: 0 auto, 2 force 3xTF32, 3 force 1xTF32; bias / act: folded into the epilogue (the caller still runs the mask pass)
  if (k < 1 || m > (1 << 30) || n > (1 << 30) || k > (1 << 30) || batch > 65535) return 0;
  // the hi/lo plane workspace is one buffer: only the main stream may use it (a side branch of a compiled
  // step takes the SIMT / fused kernels instead)
  if (S().branch != 0) return 0;
  const State::FreshPlanes fp = S().fresh;                 // planes the producer of A wrote (valid until the next launch)
  S().fresh.ptr = nullptr;
  const int emit = S().emit_next;                          // wf2cu_emit_operand_planes: this product's result feeds the next one (one shot)
  S().emit_next = 0;
  if (mode == 0) {
    // the tensor path pays two extra passes over the operands and wants >= 1 wave of 128x256 tiles
    if (m < 64 || n < 64 || k < 32) return 0;
    // below ~1 GFLOP the cluster split-K fp32 kernel (gemm.cu) wins: 21 us vs 35 us on (1024 256 784),
    // 15 us vs 26 us on 512^3; above it the tensor path does: 44 us vs 87 us on 1024^3 (gpurun_out/gemm_small2.log)
    if (2.0 * (double)m * (double)n * (double)k * (double)batch < 1.0e9) return 0;
  }
  if (!tensor_map_encoder()) return -WF2CU_ERR_CUDA;
  if (!g_attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return -cuda_fail(e, "cudaFuncSetAttribute(gemm_tc_kernel)", __FILE__, __LINE__);
    g_attr_set = true;
  }
  const int Kp = (int)((k + TK - 1) / TK) * TK;
  // 1: TF32 only (mode 3, not parity grade); 2: TF32 + bf16 corrections; 3: 3xTF32.  Modes 5 / 6 pick the tile, not the split.
  const int passes = mode == 3 ? 1 : (gemm_tc_correction() ? 3 : 2);
  const int a_nb = (sa == 0 || batch == 1) ? 1 : (int)batch;
  const int b_nb = (sb == 0 || batch == 1) ? 1 : (int)batch;
  auto plane = [&](int64_t rows, int nb) { return (((size_t)nb * rows * Kp * 4) + 255) & ~(size_t)255; };
  const size_t pa = plane(m, a_nb), pb = plane(n, b_nb);
  int rc = 0;
  CUtensorMap mah, mal, mbh, mbl, mc;
  const int c_tma = make_map_c(&mc, C, m, n, ldc, sc, batch);
  // 128x256, 256x256 or 256x192 tiles: results are bit-identical (same k order per output), pick_tile256 chooses
  const int64_t tiles_m256 = (m + T2M - 1) / T2M;
  const int64_t tiles256 = tiles_m256 * ((n + 255) / 256) * batch;
  const int64_t tiles192 = tiles_m256 * ((n + 191) / 192) * batch;
  int nb = 0;                                              // 0: the 128x256 kernel below
  bool pair2 = mode == 8;
  if (mode == 5) nb = 256;
  else if (mode == 6) nb = 192;
  else if (mode == 0 && g_t256_auto) {
    const int64_t tiles_pair = pair_override() == 0 ? 0 : ((m + 511) / 512) * ((n + 255) / 256) * batch;
    nb = pick_tile256(Kp, ((m + TM - 1) / TM) * ((n + TN - 1) / TN) * batch, tiles256, tiles192, tiles_pair, S().sm_count, passes == 2);
    if (pair_override() == 1 && m >= 512 && n >= 256) nb = 512;
  }
  if (nb == 512) pair2 = true;
  if (pair2) nb = 256;                                     // (the pair kernel shares the 256-row kernels' k-block and plane layout)
  // the second plane: fp32 lo (3xTF32) or the mixed bf16 plane laid out for the chosen kernel's k-block
  const int bf_blk = passes == 2 ? T2K : 0;                // one mixed-plane layout (k-blocks of 16) for every tile shape
  // Operand planes.  A: logical (m x k); ta = stored transposed (k x m) -> m is the contiguous axis of the source.
  // B: planes are (n x k); op(B) is k x n; !tb = stored (k x n) row-major -> n contiguous.
  //   raw_hi   a K-major, 16-byte aligned operand is its own hi plane (TMA map over the source, hi = trunc_tf32(x)): its pre-pass
  //            writes the second plane only (4 instead of 8 bytes per element)
  //   constant an operand inside a range declared with wf2cu_gemm_const_operand keeps its planes between calls (no pre-pass)
  struct Prep { const float* src; int64_t ld, sx; int rows, nb; bool rcontig, raw, konst; float* hi; float* lo; size_t pbytes; };
  Prep pr[2] = {{A, lda, sa, (int)m, a_nb, ta != 0, false, false, nullptr, nullptr, pa},
                {B, ldb, sb, (int)n, b_nb, tb == 0, false, false, nullptr, nullptr, pb}};
  size_t ws_need = 0;
  bool fresh[2] = {true, true};
  for (Prep& q : pr) {
    q.raw = passes >= 2 && !q.rcontig && rawhi_enabled() && raw_operand_ok(q.src, q.ld, q.sx, q.nb);
    q.konst = in_const_range(q.src);
    if (!q.konst) ws_need += (q.raw ? 1 : 2) * q.pbytes;
  }
  // A's second plane already written by its producer: same bits as the pre-pass would write (plane_bits == split_bf<true>)
  const bool a_fresh = fp.ptr == A && fp.lo && !ta && passes == 2 && pr[0].raw && !pr[0].konst && a_nb == 1 && fp.rows == m && fp.k == k &&
                       fp.ld == lda && Kp == k;
  if (a_fresh) ws_need -= pr[0].pbytes;
  if (ws_need && (rc = ensure_ws(ws_need))) return rc;
  {
    uint8_t* w = reinterpret_cast<uint8_t*>(g_ws);
    for (int i = 0; i < 2; ++i) {
      Prep& q = pr[i];
      if (q.konst) {
        const PlaneKey key{q.src, q.ld, q.sx, q.rows, (int)k, Kp, q.nb, q.rcontig ? 1 : 0, bf_blk, q.raw ? 1 : 0};
        const int64_t inner = q.rcontig ? q.rows : k, outer = q.rcontig ? k : q.rows;
        const size_t src_floats = (size_t)((q.nb - 1) * q.sx + (outer - 1) * q.ld + inner);
        PlaneVal* v = nullptr;
        if ((rc = const_planes(key, q.pbytes, src_floats, &v, &fresh[i]))) return rc;
        q.hi = v->hi; q.lo = v->lo;
      } else if (i == 0 && a_fresh) {
        q.lo = fp.lo; fresh[0] = false;
      } else {
        if (!q.raw) { q.hi = reinterpret_cast<float*>(w); w += q.pbytes; }
        q.lo = reinterpret_cast<float*>(w); w += q.pbytes;
      }
      if (fresh[i] && (rc = launch_split(q.src, q.hi, q.lo, q.ld, q.sx, q.rows, (int)k, Kp, q.nb, q.rcontig, bf_blk, q.raw))) return rc;
    }
  }
  // the result's own operand plane, written by the epilogue of the 256-row kernels (dense C, one batch, n a multiple of 32; declined otherwise)
  float* eplane = nullptr;
  CUtensorMap mp;
  if (emit && (pair2 || nb) && c_tma && batch == 1 && ldc == n && (n & 31) == 0) {
    eplane = planes_emit_begin(m, n);
    if (eplane && !make_map_c(&mp, eplane, m, n, n, 0, 1)) eplane = nullptr;
  }
  // hi maps: the source itself (raw) or the hi plane; lo maps: the second plane
  auto maps = [&](const Prep& q, CUtensorMap* mh, CUtensorMap* ml, int box_k, int box_rows) -> int {
    int r = q.raw ? make_map_k(mh, q.src, k, q.rows, q.ld, q.nb, q.sx, box_k, box_rows)
                  : make_map_k(mh, q.hi, Kp, q.rows, Kp, q.nb, (int64_t)q.rows * Kp, box_k, box_rows);
    if (r) return r;
    return make_map_k(ml, q.lo, Kp, q.rows, Kp, q.nb, (int64_t)q.rows * Kp, box_k, box_rows);
  };

  if (pair2) {
    // CTA pair, 512 x 256 pair tiles: each CTA loads its 256 rows of A and HALF of the B tile (gemm_tc2sm.cuh)
    static bool attr2 = false;
    if (!attr2) {
      cudaError_t e = cudaFuncSetAttribute(gemm_tc2sm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, P2_SMEM);
      if (e != cudaSuccess) return -cuda_fail(e, "cudaFuncSetAttribute(gemm_tc2sm_kernel)", __FILE__, __LINE__);
      attr2 = true;
    }
    if ((rc = maps(pr[0], &mah, &mal, T2K, 256))) return rc;
    if ((rc = maps(pr[1], &mbh, &mbl, T2K, 128))) return rc;
    TcArgs t;
    t.C = C; t.ldc = ldc; t.sc = sc; t.m = (int)m; t.n = (int)n; t.nkb = Kp / T2K;
    t.tiles_m = (int)((m + 511) / 512); t.tiles_n = (int)((n + 255) / 256); t.batch = (int)batch;
    t.passes = passes;
    t.c_vec = ((reinterpret_cast<uintptr_t>(C) & 15) == 0 && (ldc & 3) == 0 && (sc & 3) == 0) ? 1 : 0;
    t.a_bcast = a_nb == 1 ? 1 : 0;
    t.b_bcast = b_nb == 1 ? 1 : 0;
    t.bias = bias; t.act = act; t.res = res; t.ldres = ldres; t.c_tma = c_tma; t.plane = eplane ? 1 : 0; if (eplane) t.pmap = mp;
    const int64_t total = (int64_t)t.tiles_m * t.tiles_n * t.batch;
    {
      // raster: 8 m-tiles of 512 rows per group — with 74 pairs in flight that is ~8 x 9 pair tiles sharing A blocks 9-fold and B
      // blocks 8-fold (16 per group left 4.6-fold A reuse: 5.9 GB from DRAM at 8192^3 against 3.3 GB for the single-CTA kernel)
      static int grp = -1;
      if (grp < 0) { const char* e = getenv("WF2CU_PAIR_GROUP"); grp = e ? atoi(e) : 8; if (grp < 1) grp = 8; }
      t.group_m = grp;
    }
    const int64_t pairs_max = S().sm_count / 2;
    const int grid = 2 * (int)(total < pairs_max ? total : pairs_max);
    gemm_tc2sm_kernel<<<grid, P2_THREADS, P2_SMEM, S().stream>>>(mah, mal, mbh, mbl, mc, t);
    rc = post_launch("wf2cu_gemm(tcgen05 CTA pair)");
    if (!rc && eplane) planes_emit_publish(C, m, n, ldc, eplane);
    S().gemm_last_path = WF2CU_GEMM_PATH_TC_2SM;
    return rc ? -rc : 1;
  }
  if (nb) {
    // 256-row tiles: two M = 128 MMAs share every B slice (42.7 instead of 62.5 B/clk of L2 -> SM traffic per SM)
    static bool attr256 = false;
    if (!attr256) {
      cudaError_t e = cudaFuncSetAttribute(gemm_tc256_kernel<256, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, t2_smem_bytes(256, 3));
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(gemm_tc256_kernel<192, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, t2_smem_bytes(192, 3));
      if (e != cudaSuccess) return -cuda_fail(e, "cudaFuncSetAttribute(gemm_tc256_kernel)", __FILE__, __LINE__);
      attr256 = true;
    }
    if ((rc = maps(pr[0], &mah, &mal, T2K, 256))) return rc;
    if ((rc = maps(pr[1], &mbh, &mbl, T2K, nb))) return rc;
    TcArgs t;
    t.C = C; t.ldc = ldc; t.sc = sc; t.m = (int)m; t.n = (int)n; t.nkb = Kp / T2K;
    t.tiles_m = (int)tiles_m256; t.tiles_n = (int)((n + nb - 1) / nb); t.batch = (int)batch;
    t.passes = passes;
    t.c_vec = ((reinterpret_cast<uintptr_t>(C) & 15) == 0 && (ldc & 3) == 0 && (sc & 3) == 0) ? 1 : 0;
    t.a_bcast = a_nb == 1 ? 1 : 0;
    t.b_bcast = b_nb == 1 ? 1 : 0;
    t.bias = bias; t.act = act; t.res = res; t.ldres = ldres; t.c_tma = c_tma; t.plane = eplane ? 1 : 0; if (eplane) t.pmap = mp;
    const int64_t total = (int64_t)t.tiles_m * t.tiles_n * t.batch;
    const int grid = (int)(total < S().sm_count ? total : S().sm_count);
    if (nb == 256) gemm_tc256_kernel<256, 3><<<grid, EPI8_THREADS, t2_smem_bytes(256, 3), S().stream>>>(mah, mal, mbh, mbl, mc, t);
    else           gemm_tc256_kernel<192, 3><<<grid, EPI8_THREADS, t2_smem_bytes(192, 3), S().stream>>>(mah, mal, mbh, mbl, mc, t);
    rc = post_launch(nb == 256 ? "wf2cu_gemm(tcgen05 256x256)" : "wf2cu_gemm(tcgen05 256x192)");
    if (!rc && eplane) planes_emit_publish(C, m, n, ldc, eplane);
    S().gemm_last_path = nb == 256 ? WF2CU_GEMM_PATH_TC_256X256 : WF2CU_GEMM_PATH_TC_256X192;
    return rc ? -rc : 1;
  }
  if ((rc = maps(pr[0], &mah, &mal, TK, TM))) return rc;
  if ((rc = maps(pr[1], &mbh, &mbl, TK, TN))) return rc;

  TcArgs t;
  t.C = C; t.ldc = ldc; t.sc = sc; t.m = (int)m; t.n = (int)n; t.nkb = Kp / TK;
  t.tiles_m = (int)((m + TM - 1) / TM); t.tiles_n = (int)((n + TN - 1) / TN); t.batch = (int)batch;
  t.passes = passes;
  t.c_vec = ((reinterpret_cast<uintptr_t>(C) & 15) == 0 && (ldc & 3) == 0 && (sc & 3) == 0) ? 1 : 0;
  t.a_bcast = a_nb == 1 ? 1 : 0;
  t.b_bcast = b_nb == 1 ? 1 : 0;
  t.bias = bias; t.act = act; t.res = res; t.ldres = ldres; t.c_tma = c_tma;
  const int64_t total = (int64_t)t.tiles_m * t.tiles_n * t.batch;
  const int grid = (int)(total < S().sm_count ? total : S().sm_count);
  gemm_tc_kernel<<<grid, GEMM_THREADS, SMEM_BYTES, S().stream>>>(mah, mal, mbh, mbl, mc, t);
  rc = post_launch("wf2cu_gemm(tcgen05)");
  S().gemm_last_path = WF2CU_GEMM_PATH_TC_128X256;
  return rc ? -rc : 1;
}

// ---- raw-operand kernel (gemm_raw.cuh): A must be K-major (not transposed); B raw when K-major (trans_b), else planes
static int make_map_raw(CUtensorMap* map, const float* base, int64_t k, int64_t rows, int64_t ld, int64_t nbatch, int64_t bstride, int box_rows) {
  cuuint64_t dims[3] = {(cuuint64_t)k, (cuuint64_t)rows, (cuuint64_t)(nbatch < 1 ? 1 : nbatch)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 4, (cuuint64_t)(nbatch > 1 ? bstride : rows * ld) * 4};
  cuuint32_t box[3] = {(cuuint32_t)RK, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return -set_error(WF2CU_ERR_CUDA, "cuTensorMapEncodeTiled(raw operand) failed (%d) k=%lld rows=%lld ld=%lld", (int)r,
                                           (long long)k, (long long)rows, (long long)ld);
  return 0;
}

static bool raw_operand_ok(const float* p, int64_t ld, int64_t bstride, int64_t nbatch) {
  return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && (ld & 3) == 0 && (nbatch <= 1 || ((bstride & 3) == 0 && bstride > 0));
}

// Auto rule: raw-operand kernel or planes kernels?  Cycle model fitted on B200 (scripts/gemm_raw_bench.py):
// per k-block of 16 a tile costs 1180 cycles on the raw kernel (shared-memory bound: TMA writes + converter traffic + operand reads
// of six 128x256x8 MMAs), 1110 / 1850 / 1500 on the 128x256 / 256x256 / 256x192 planes kernels; the 256-row tiles expose their
// epilogue; an operand pre-pass moves 12 bytes per element at ~4.5 TB/s plus a launch.  The raw kernel skips A's pre-pass always
// and B's when B is K-major.  2048^3: raw 88 us vs 101; 4096^3: 654 vs 571; 8192x2304x768: 124 vs 132; 8192x768x3072: 236 vs 210.
// share of a full pre-pass (read 4 + write 8 bytes per element, one launch) an operand of the planes kernels costs: 0 when its
// planes are kept (constant operand), 2/3 when the source is its own hi plane, 1 otherwise
double gemm_tc_prepass_share(const float* p, int64_t ld, int64_t bstride, int64_t nbatch, bool kmajor) {
  if (in_const_range(p)) return 0.0;
  if (kmajor && S().fresh.ptr == p && S().fresh.lo && S().fresh.ld == ld) return 0.0;   // its producer wrote the plane
  if (kmajor && rawhi_enabled() && gemm_tc_correction() == 0 && raw_operand_ok(p, ld, bstride, nbatch)) return 2.0 / 3.0;
  return 1.0;
}

bool gemm_raw_preferred(int tb, int64_t m, int64_t n, int64_t k, int64_t batch, double a_share, double b_share) {
  const double sms = (double)S().sm_count;
  auto waves = [&](int64_t tiles) { return (double)((tiles * batch + (int64_t)sms - 1) / (int64_t)sms); };
  auto presplit = [&](int64_t rows) { return (double)rows * (double)k * (double)batch * 12.0 / 4.5e12 * 1.9e9 + 6000.0; };
  const double nkb = (double)((k + 15) / 16);
  const int64_t t128 = ((m + 127) / 128) * ((n + 255) / 256), t256 = ((m + 255) / 256) * ((n + 255) / 256), t192 = ((m + 255) / 256) * ((n + 191) / 192);
  const double raw = waves(t128) * nkb * 1180.0 + 8000.0 + (tb ? 0.0 : (b_share > 0.0 ? presplit(n) : 0.0));
  // the planes kernels with bf16 corrections (the default) spend 2 tensor-pipe units per k instead of 3: 870 / 1470 / 1190
  // (r02d: 2048^3 raw 90 us vs 95; 8192x2304x768 raw 129 vs 112; 4096^3 raw 654 vs 490)
  const bool bf = gemm_tc_correction() == 0;
  double planes = waves(t128) * nkb * (bf ? 870.0 : 1110.0);
  const double p256 = waves(t256) * (nkb * (bf ? 1470.0 : 1850.0) + (bf ? 7000.0 : 10000.0));
  const double p192 = waves(t192) * (nkb * (bf ? 1190.0 : 1500.0) + (bf ? 5500.0 : 8000.0));
  if (p256 < planes) planes = p256;
  if (p192 < planes) planes = p192;
  if (pair_override() != 0) {          // CTA-pair kernel: ~6 % fewer cycles per k-block than the 256x256 tiles, on sms / 2 pairs
    const int64_t tp = ((m + 511) / 512) * ((n + 255) / 256) * batch, pairs = (int64_t)sms / 2;
    const double ppair = (double)((tp + pairs - 1) / pairs) * (nkb * (bf ? 1385.0 : 1755.0) + (bf ? 7000.0 : 10000.0));
    if (ppair < planes) planes = ppair;
  }
  planes += (a_share > 0.0 ? a_share * presplit(m) : 0.0) + (b_share > 0.0 ? b_share * presplit(n) : 0.0);
  return raw <= planes;
}

// returns 1 when it handled the problem, 0 when it declines, <0 = -(error code)
int gemm_raw_try(int ta, int tb, int64_t m, int64_t n, int64_t k, const float* A, int64_t lda, const float* B, int64_t ldb,
                 float* C, int64_t ldc, int64_t batch, int64_t sa, int64_t sb, int64_t sc, const float* bias, int act, const float* res,
                 int64_t ldres) {
  if (ta || k < 1 || m > (1 << 30) || n > (1 << 30) || k > (1 << 30) || batch > 65535 || S().branch != 0) return 0;
  const int a_nb = (sa == 0 || batch == 1) ? 1 : (int)batch;
  const int b_nb = (sb == 0 || batch == 1) ? 1 : (int)batch;
  if (!raw_operand_ok(A, lda, sa, a_nb)) return 0;
  const bool b_raw = tb != 0 && raw_operand_ok(B, ldb, sb, b_nb);
  if (!tensor_map_encoder()) return -WF2CU_ERR_CUDA;
  static bool attr_set = false;
  if (!attr_set) {
    const int bytes = raw_smem(RawGemmJob::kEpiWarps, RawGemmJob::kEpiBytes);
    cudaError_t e = cudaFuncSetAttribute(gemm_raw_kernel<RawGemmJob, true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_raw_kernel<RawGemmJob, true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return -cuda_fail(e, "cudaFuncSetAttribute(gemm_raw_kernel)", __FILE__, __LINE__);
    attr_set = true;
  }
  CUtensorMap ma, mb, mbl, mc;
  memset(&mbl, 0, sizeof(mbl));
  int rc;
  if ((rc = make_map_raw(&ma, A, k, m, lda, a_nb, sa, RM))) return rc;
  if (b_raw) {
    if ((rc = make_map_raw(&mb, B, k, n, ldb, b_nb, sb, RN))) return rc;
    mbl = mb;
  } else {
    // B is MN-major for the product (or misaligned): transposing split into K-major hi / lo planes, as before
    const int Kp = (int)((k + TK - 1) / TK) * TK;
    const size_t pb = ((((size_t)b_nb * n * Kp * 4) + 255) & ~(size_t)255);
    float *bh, *bl;
    bool fresh = true;
    if (in_const_range(B)) {              // a constant operand keeps its planes between calls
      const bool rc_b = tb == 0;
      const PlaneKey key{B, ldb, sb, (int)n, (int)k, Kp, b_nb, rc_b ? 1 : 0, 0, 0};
      const int64_t inner = rc_b ? n : k, outer = rc_b ? k : n;
      PlaneVal* v = nullptr;
      if ((rc = const_planes(key, pb, (size_t)((b_nb - 1) * sb + (outer - 1) * ldb + inner), &v, &fresh))) return rc;
      bh = v->hi; bl = v->lo;
    } else {
      if ((rc = ensure_ws(2 * pb))) return rc;
      bh = g_ws;
      bl = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bh) + pb);
    }
    if (fresh && (rc = launch_split(B, bh, bl, ldb, sb, (int)n, (int)k, Kp, b_nb, tb == 0, 0))) return rc;
    if ((rc = make_map256(&mb, bh, (int)n, Kp, b_nb, RN))) return rc;
    if ((rc = make_map256(&mbl, bl, (int)n, Kp, b_nb, RN))) return rc;
  }
  RawGemmJob job;
  TcArgs& t = job.t;
  t.C = C; t.ldc = ldc; t.sc = sc; t.m = (int)m; t.n = (int)n; t.nkb = (int)((k + RK - 1) / RK);
  t.tiles_m = (int)((m + RM - 1) / RM); t.tiles_n = (int)((n + RN - 1) / RN); t.batch = (int)batch;
  // 3xTF32: the converter warps are this kernel's bottleneck, and making mixed bf16 tiles (two scattered 8-byte stores per chunk
  // instead of one 16-byte store) costs them more than the two saved MMAs give back (r02f: 4096^3 654 -> 748 us)
  t.passes = 3;
  t.c_vec = ((reinterpret_cast<uintptr_t>(C) & 15) == 0 && (ldc & 3) == 0 && (sc & 3) == 0) ? 1 : 0;
  t.a_bcast = a_nb == 1 ? 1 : 0;
  t.b_bcast = b_nb == 1 ? 1 : 0;
  t.bias = bias; t.act = act; t.res = res; t.ldres = ldres;
  t.c_tma = make_map_c(&mc, C, m, n, ldc, sc, batch);
  const int64_t total = (int64_t)t.tiles_m * t.tiles_n * t.batch;
  const int grid = (int)(total < S().sm_count ? total : S().sm_count);
  const int threads = raw_threads(RawGemmJob::kEpiWarps, true), bytes = raw_smem(RawGemmJob::kEpiWarps, RawGemmJob::kEpiBytes);
  cudaError_t e = b_raw ? launch_pdl(gemm_raw_kernel<RawGemmJob, true, true, true>, dim3(grid), dim3(threads), bytes, S().stream, ma, ma, mb, mbl, mc, job)
                        : launch_pdl(gemm_raw_kernel<RawGemmJob, true, false, true>, dim3(grid), dim3(threads), bytes, S().stream, ma, ma, mb, mbl, mc, job);
  if (e != cudaSuccess) return -cuda_fail(e, "launch(gemm_raw_kernel)", __FILE__, __LINE__);
  rc = post_launch("wf2cu_gemm(tcgen05 raw operands)");
  S().gemm_last_path = WF2CU_GEMM_PATH_TC_RAW;
  return rc ? -rc : 1;
}

// lo planes of Q and K for the first 2E features of every (b, n) row of the QKV activation: fp32 lo[b, n, c] = x - trunc_tf32(x),
// or (mixed != 0) per 16 features bf16_rn(x[0:16]) then bf16_rn(x[0:16] - trunc_tf32(x[0:16])) in the same 64 bytes
__global__ void __launch_bounds__(256) attn_split_qk_kernel(const float* __restrict__ qkv, float* __restrict__ lo, int64_t rows, int mixed) {
  pdl_prologue();
  constexpr int V4 = 2 * RA_E / 4;                                   // float4 per row
  const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= rows * V4) return;
  const int64_t r = idx / V4;
  const int c4 = (int)(idx - r * V4);
  const float4 x = reinterpret_cast<const float4*>(qkv + r * (3 * RA_E))[c4];
  if (mixed) {
    // features [4 c4, 4 c4 + 4) of block c4 / 4: hb at 16-bit index blk * 32 + j, lb at blk * 32 + 16 + j
    uint16_t* row = reinterpret_cast<uint16_t*>(lo + r * (2 * RA_E));
    uint16_t* dst = row + (c4 >> 2) * 32 + (c4 & 3) * 4;
    *reinterpret_cast<uint2*>(dst) = make_uint2(raw_bf2(raw_fin(x.x), raw_fin(x.y)), raw_bf2(raw_fin(x.z), raw_fin(x.w)));
    *reinterpret_cast<uint2*>(dst + 16) = make_uint2(raw_bf2(raw_lo(x.x), raw_lo(x.y)), raw_bf2(raw_lo(x.z), raw_lo(x.w)));
    return;
  }
  reinterpret_cast<float4*>(lo + r * (2 * RA_E))[c4] = make_float4(raw_lo(x.x), raw_lo(x.y), raw_lo(x.z), raw_lo(x.w));
}

// The GPT-2 example's attention weights through the raw main loop: P[b,h,i,j] = exp(s_h / divisor) / sum_h' exp(s_h' / divisor).
// qkv: dense (B N 3E), 16-byte aligned; P: dense (B H N N); N a multiple of 128.  Returns 0 or a wf2cu status.
int attn_scores_raw(const float* qkv, int64_t batch, int64_t seq, float divisor, float* P) {
  if (!tensor_map_encoder()) return WF2CU_ERR_CUDA;
  if (S().branch != 0) return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_attn_heads_softmax: main stream only (the workspace is not per branch)");
  const int B = (int)batch, N = (int)seq;
  // WF2CU_ATTN_PLANES=1: lo planes of Q and K pre-split into the workspace and loaded by TMA instead of made by converter warps
  // (twice the L2 -> SM bytes: the kernel is then bound by the ~7.4 TB/s the L2 delivers to the SMs; kept for A/B measurements)
  static int planes = -1;
  if (planes < 0) { const char* e = getenv("WF2CU_ATTN_PLANES"); planes = (e && e[0] == '1') ? 1 : 0; }
  // corrections of Q.K^T as bf16 MMAs on mixed lo planes: only when the planes are pre-split (the converter warps of the raw
  // variant are its bottleneck; mixed tiles cost them more than the saved MMAs return: 194 -> 336 us, r02f)
  const int hyb = (planes && gemm_tc_correction() == 0) ? 1 : 0;
  float* qk_lo = nullptr;
  if (planes) {
    if (ensure_ws((size_t)B * N * 2 * RA_E * 4)) return WF2CU_ERR_CUDA;
    qk_lo = g_ws;
    const int64_t rows = batch * seq, work = rows * (2 * RA_E / 4);
    WF2_CUDA(launch_pdl(attn_split_qk_kernel, dim3((unsigned)((work + 255) / 256)), dim3(256), 0, S().stream, qkv, qk_lo, rows, hyb));
    const int rc0 = post_launch("wf2cu_attn_heads_softmax(split Q K)");
    if (rc0) return rc0;
  }
  static bool attr_set = false;
  const int bytes = raw_smem(RawAttnJob::kEpiWarps, RawAttnJob::kEpiBytes);
  if (!attr_set) {
    WF2_CUDA(cudaFuncSetAttribute(gemm_raw_kernel<RawAttnJob, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    WF2_CUDA(cudaFuncSetAttribute(gemm_raw_kernel<RawAttnJob, true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    attr_set = true;
  }
  CUtensorMap mq, mql, mk, mkl, mp;
  {
    cuuint32_t boxq[3] = {RK, RM, 1}, boxk[3] = {RK, RN, 1}, estr[4] = {1, 1, 1, 1};
    cuuint64_t dims[3] = {(cuuint64_t)(3 * RA_E), (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)(3 * RA_E) * 4, (cuuint64_t)N * 3 * RA_E * 4};
    cuuint64_t ldims[3] = {(cuuint64_t)(2 * RA_E), (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t lstrides[2] = {(cuuint64_t)(2 * RA_E) * 4, (cuuint64_t)N * 2 * RA_E * 4};
    cuuint64_t pd[4] = {(cuuint64_t)N, (cuuint64_t)N, (cuuint64_t)RA_H, (cuuint64_t)B};
    cuuint64_t ps[3] = {(cuuint64_t)N * 4, (cuuint64_t)N * N * 4, (cuuint64_t)RA_H * N * N * 4};
    cuuint32_t pb[4] = {8, 32, 1, 1};
    CUresult r[5];
    r[0] = g_encode(&mq, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)qkv, dims, strides, boxq, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    r[1] = g_encode(&mk, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)qkv, dims, strides, boxk, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    r[2] = r[3] = CUDA_SUCCESS;
    mql = mq; mkl = mk;
    if (planes) {
      r[2] = g_encode(&mql, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)qk_lo, ldims, lstrides, boxq, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      r[3] = g_encode(&mkl, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)qk_lo, ldims, lstrides, boxk, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    r[4] = g_encode(&mp, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)P, pd, ps, pb, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    for (int i = 0; i < 5; ++i)
      if (r[i] != CUDA_SUCCESS) return set_error(WF2CU_ERR_CUDA, "cuTensorMapEncodeTiled(attention scores, map %d) failed (%d)", i, (int)r[i]);
  }
  static int dbg = -1;
  if (dbg < 0) { const char* e = getenv("WF2CU_ATTN_DBG"); dbg = e ? atoi(e) : 0; }
  RawAttnJob job{RawAttnArgs{B, N, (N + RM - 1) / RM, (N + RN - 1) / RN, (float)(1.4426950408889634 / (double)divisor), dbg, hyb}};
  const int items = job.a.B * job.a.tiles_i * job.a.tiles_j, sms = S().sm_count;
  const int grid = items < sms ? items : sms;
  if (planes) WF2_CUDA(launch_pdl(gemm_raw_kernel<RawAttnJob, false, false, true>, dim3(grid), dim3(raw_threads(RawAttnJob::kEpiWarps, false)), bytes,
                                  S().stream, mq, mql, mk, mkl, mp, job));
  else        WF2_CUDA(launch_pdl(gemm_raw_kernel<RawAttnJob, true, true, true>, dim3(grid), dim3(raw_threads(RawAttnJob::kEpiWarps, true)), bytes,
                                  S().stream, mq, mql, mk, mkl, mp, job));
  return post_launch("wf2cu_attn_heads_softmax");
}

// One-pass form of the above for wf2cu_attn_fused: E[b,h,i,j] = exp(s_h / divisor) (un-normalised) and R[b,i,j] = RN(1 / sum_h E).
// E: dense (B H N N), R: dense (B N N), both 16-byte aligned.  Returns 0 or a wf2cu status.
int attn_scores_onepass(const float* qkv, int64_t batch, int64_t seq, float divisor, float* E, float* R) {
  if (!tensor_map_encoder()) return WF2CU_ERR_CUDA;
  const int B = (int)batch, N = (int)seq;
  static bool attr_set = false;
  const int bytes = raw_smem(RawAttnJob1::kEpiWarps, RawAttnJob1::kEpiBytes);
  if (!attr_set) {
    WF2_CUDA(cudaFuncSetAttribute(gemm_raw_kernel<RawAttnJob1, true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    attr_set = true;
  }
  CUtensorMap mq, mk, me, mr;
  {
    cuuint32_t boxq[3] = {RK, RM, 1}, boxk[3] = {RK, RN, 1}, estr[4] = {1, 1, 1, 1};
    cuuint64_t dims[3] = {(cuuint64_t)(3 * RA_E), (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)(3 * RA_E) * 4, (cuuint64_t)N * 3 * RA_E * 4};
    cuuint64_t pd[4] = {(cuuint64_t)N, (cuuint64_t)N, (cuuint64_t)RA_H, (cuuint64_t)B};
    cuuint64_t ps[3] = {(cuuint64_t)N * 4, (cuuint64_t)N * N * 4, (cuuint64_t)RA_H * N * N * 4};
    cuuint32_t pb[4] = {8, 32, 1, 1};
    cuuint64_t rd[3] = {(cuuint64_t)N, (cuuint64_t)N, (cuuint64_t)B};
    cuuint64_t rs[2] = {(cuuint64_t)N * 4, (cuuint64_t)N * N * 4};
    cuuint32_t rb[3] = {8, 32, 1};
    CUresult r[4];
    r[0] = g_encode(&mq, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)qkv, dims, strides, boxq, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    r[1] = g_encode(&mk, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)qkv, dims, strides, boxk, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    r[2] = g_encode(&me, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)E, pd, ps, pb, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    r[3] = g_encode(&mr, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)R, rd, rs, rb, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    for (int i = 0; i < 4; ++i)
      if (r[i] != CUDA_SUCCESS) return set_error(WF2CU_ERR_CUDA, "cuTensorMapEncodeTiled(attention, one pass, map %d) failed (%d)", i, (int)r[i]);
  }
  RawAttnJob1 job{RawAttnArgs{B, N, (N + RM - 1) / RM, (N + RN - 1) / RN, (float)(1.4426950408889634 / (double)divisor), 0, 0}};
  const int items = job.a.B * job.a.tiles_i * job.a.tiles_j, sms = S().sm_count;
  const int grid = items < sms ? items : sms;
  // map slots: A = Q, A-lo slot = R (store), B = K, B-lo slot = K (unused), C = E (store)
  WF2_CUDA(launch_pdl(gemm_raw_kernel<RawAttnJob1, true, true, true>, dim3(grid), dim3(raw_threads(RawAttnJob1::kEpiWarps, true)), bytes,
                      S().stream, mq, mr, mk, mk, me, job));
  return post_launch("wf2cu_attn_fused(scores)");
}

// returns 1 when it handled the problem, 0 when it declines, <0 = -(error code)
int gemm_tc_fused_try(int ta, int tb, int64_t m, int64_t n, int64_t k, const float* A, int64_t lda, const float* B, int64_t ldb,
                      float* C, int64_t ldc, int64_t batch, int64_t sa, int64_t sb, int64_t sc, const float* bias, int act,
                      const float* mask, int64_t ldmask) {
  if (m < 1 || n < 1 || k < 1 || m > (1 << 24) || n > (1 << 24) || k > (1 << 24) || batch < 1 || batch > 8000) return 0;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM);
    if (e != cudaSuccess) return -cuda_fail(e, "cudaFuncSetAttribute(gemm_tc_fused_kernel)", __FILE__, __LINE__);
    attr_set = true;
  }
  const int64_t tiles = ((m + FM - 1) / FM) * ((n + FN - 1) / FN) * batch;
  if ((m + FM - 1) / FM > 65535) return 0;
  const int nkb_total = (int)((k + TK - 1) / TK);
  int ksplit = 1;
  // spread K over a cluster while the grid is below ~2 CTAs per SM and every slice keeps >= 2 k-blocks
  while (ksplit < 8 && tiles * ksplit * 2 <= (int64_t)S().sm_count * 2 && nkb_total / (ksplit * 2) >= 2) ksplit *= 2;
  const int kchunk = ((nkb_total + ksplit - 1) / ksplit) * TK;
  FusedGemmArgs g;
  memset(&g, 0, sizeof(g));
  g.A = A; g.B = B; g.C = C; g.lda = lda; g.ldb = ldb; g.ldc = ldc;
  g.m = (int)m; g.n = (int)n; g.k = (int)k; g.ta = ta; g.tb = tb;
  g.sa = sa; g.sb = sb; g.sc = sc;
  g.a_vec = ((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (lda & 3) == 0 && (sa & 3) == 0) ? 1 : 0;
  g.b_vec = ((reinterpret_cast<uintptr_t>(B) & 15) == 0 && (ldb & 3) == 0 && (sb & 3) == 0) ? 1 : 0;
  g.c_vec = ((reinterpret_cast<uintptr_t>(C) & 15) == 0 && (ldc & 3) == 0 && (sc & 3) == 0) ? 1 : 0;
  g.ksplit = ksplit; g.kchunk = kchunk;
  g.bias = bias; g.mask = mask; g.ldmask = ldmask; g.act = act;
  {
    static int hint = -1;
    // measured on GPT-2's P.V (B200): 221 / 212 / 207 / 213 us for hints 0 / 1 / 2 / 3 (profiles/r01d_attention_ncu.md)
    if (hint < 0) { const char* e = getenv("WF2CU_FUSED_AHINT"); hint = e ? atoi(e) : 2; if (hint < 0 || hint > 3) hint = 2; }
    if (!ta && kchunk / TK >= 16) g.a_hint = hint;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)((n + FN - 1) / FN), (unsigned)((m + FM - 1) / FM), (unsigned)(batch * ksplit));
  cfg.blockDim = dim3(256, 1, 1);
  cfg.dynamicSmemBytes = F_SMEM;
  cfg.stream = S().stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = (unsigned)ksplit;
  cfg.attrs = attr; cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_tc_fused_kernel, g);
  if (e != cudaSuccess) return -cuda_fail(e, "cudaLaunchKernelEx(gemm_tc_fused_kernel)", __FILE__, __LINE__);
  const int rc = post_launch("wf2cu_gemm(tcgen05 fused)");
  S().gemm_last_path = WF2CU_GEMM_PATH_TC_FUSED;
  return rc ? -rc : 1;
}

}  // namespace wf2
