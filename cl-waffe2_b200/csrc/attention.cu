// attention.cu — the GPT-2 example's SelfAttention as two tcgen05 kernels (second half).
//
// The reference composes it from nodes (`examples/gpt-2/utils.lisp:204-244`): split-heads (view + permute +
// ->contiguous copies), `(!matmul Q K)` into a materialised (B H N N) tensor, `(!div w N)` with N = the LAST axis of w
// (the sequence length, not sqrt(d)), `(!softmax w :avoid-overflow nil)` with !softmax's default axis 1 = the HEADS axis
// of the 4-D tensor, no causal mask, `(!matmul w V)`, merge-heads (permute + reshape copy).  All of it is reproduced as
// written; what changes is the number of passes over the 402 MB (B H N N) tensor at B = 8, N = 1024:
//
//   wf2cu_attn_heads_softmax : P[b,h,i,j] = e[b,h,i,j] / sum_h' e[b,h',i,j],  e = exp((Q_h[i,:] . K_h[j,:]) / N)
//       reads Q and K straight from the dense (B N 3E) activation of the QKV projection (both are K-major there: no
//       head-split copies, no transposes, no hi / lo planes in global memory) and writes P ONCE.  It runs on the raw-operand
//       3xTF32 main loop of MatmulNode (gemm_raw.cuh: 128 queries x 256 keys per tile, 128x256x8 MMAs) with two epilogues:
//         pass 1  den[i,j] = sum over the 12 heads (in head order, like `!sum`) of exp(s / N), accumulated in registers across
//                 the 12 head sub-tiles of a (128 queries x 256 keys) position — it never leaves the SM
//         pass 2  the product again (Q.K is 13 GFLOP per layer: cheap next to a second trip of the 402 MB tensor through HBM),
//                 p = exp(s / N) * RN(1 / den), 32 x 8 blocks through the bulk-copy engine into P
//       How it got there (all measured on B200, B = 8, N = 1024, per layer):  the heads-axis softmax couples the 12 heads of
//       one (i, j), so the first design kept all 12 score tiles of a (128 x 32) tile in TMEM: 144 us with ONE truncated TF32
//       pass (error 3e-4 on P for N(0, 2) inputs: not acceptable for a library kernel), 271 us with the compensated three —
//       whether the lo planes came from global memory, from converter warps, or the Q operand lived in tensor memory (294 us).
//       scripts/micro/mma_rate_test.cu (first version; its 45-cycle floor was the ELECT loop of the issue path): a kind::tf32 MMA with M = 128 costs max(45, N / 2) cycles, so the 288 N = 32 MMAs of a
//       tile ran at a third of the tensor rate and the 16-warp exp/divide epilogue could not hide behind them.  With N = 256
//       per MMA the products of BOTH passes together take 147 us of main loop; the whole kernel runs in 194 us.
//   wf2cu_attn_pv : out[b,i,h*64:(h+1)*64] = sum_j P[b,h,i,j] V_h[j,:], error-compensated 3xTF32 (as MatmulNode), written
//       in the merged-heads layout (B N E) that merge-heads produces (no permute copies).  P tiles arrive by TMA as RAW
//       fp32; converter warps move each tile into tensor memory as the A operand (raw bits = the TF32 hi plane, lo = x -
//       trunc(x)); V's hi / lo planes are split and transposed once per call (25 MB against 402 MB of P).  warp 0 TMA,
//       warp 1 MMA (12 x 128x64x8 per k-block into one of two 64-column accumulators), warps 2-5 converters, warps 6-9 epilogue.
//   wf2cu_attn_fused : both in one call, the heads visited ONCE: the score kernel (gemm_raw.cuh, RawAttnJob1) writes the un-normalised
//       exponentials E and the reciprocal denominators R = 1 / sum_h E; attn_pv6_kernel (six heads per CTA, R held in registers across
//       them) forms p = RN(e r) on the way into tensor memory.  The V split runs BESIDE the score kernel (programmatic dependent launch:
//       see attn_split_v_kernel).  Bit-identical to the two launches.
// All kernels are specialised for 12 heads of 64 features (GPT-2 small); other shapes return WF2CU_ERR_UNSUPPORTED and the
// host issues the nodes one by one, exactly as before.
#include <cuda.h>
#include "common.cuh"
#include "math.cuh"
#include "tc.cuh"

namespace wf2 {

constexpr int AH = 12, AD = 64, AE = AH * AD;
constexpr int UK = 8;

// ------------------------------------------------------------------ helpers
__device__ __forceinline__ void tc_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
               "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
               ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
                 "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]),
                 "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]),
                 "r"(v[30]), "r"(v[31]) : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem desc], kind::tf32: row r of the 128-row A tile is TMEM lane r, element k of it column a + k
// (scripts/micro/ts_mma_test.cu checks exactly this layout on hardware)
__device__ __forceinline__ void tc_mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n"
               ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ float tf32_lo(float x) {   // x - trunc_tf32(x), exact; inf / nan live in the hi part only
  return (fabsf(x) <= 3.0e38f) ? x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u) : 0.f;
}

// ------------------------------------------------------------------ P . V
constexpr int P_M = 128, P_N = 64, P_K = 32;
constexpr int P_ATILE = P_M * P_K * 4;               // 16 KiB
constexpr int P_BTILE = P_N * P_K * 4;               // 8 KiB
constexpr int P_STAGE = P_ATILE + 2 * P_BTILE;       // raw P | V_hi | V_lo = 32 KiB
constexpr int P_NS = 6;
// NORM variant (wf2cu_attn_fused): the P tile holds the UN-normalised exponentials and a fourth tile of the stage the reciprocal
// denominators r[i, j] of the same positions (shared by the 12 heads): stage = E | V_hi | V_lo | R = 48 KiB, 4 stages
__host__ __device__ constexpr int p_stage(bool norm) { return P_STAGE + (norm ? P_ATILE : 0); }
__host__ __device__ constexpr int p_ns(bool norm) { return norm ? 4 : P_NS; }
__host__ __device__ constexpr int p_smem(bool norm) { return p_ns(norm) * p_stage(norm) + 1024 + 256; }
constexpr int P_ACOL = 2 * P_N;                      // TMEM: columns [0, 128) two accumulators, [128, 256) two P operand buffers
constexpr int P_THREADS = 320;
constexpr int P_SMEM = P_NS * P_STAGE + 1024 + 256;
constexpr uint32_t P_IDESC = idesc_tf32_mn(P_M, P_N);

// V_h (N x 64 per (b, h), rows 3E apart inside the QKV activation) -> hi / lo planes [b*H + h][64][N], K-major
// late = 1 (wf2cu_attn_fused): the kernel launched before this one is the score kernel, which does not produce anything read here and
// lets its dependents in only after the producer of qkv is complete (gemm_raw_kernel<RawAttnJob1>).  So this grid does not wait at its
// start — it runs beside the score tiles, which leave half of the issue slots and most of the DRAM bandwidth idle — and ONE of its CTAs
// waits for the score kernel before it exits: the grid as a whole is complete only then, which is what the P.V kernel's own
// griddepcontrol.wait relies on for E and R.
__global__ void __launch_bounds__(256) attn_split_v_kernel(const float* __restrict__ qkv, float* __restrict__ hi, float* __restrict__ lo, int N, int late) {
  __shared__ float t[32][65];
  if (late) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  else pdl_prologue();
  const int bh = blockIdx.y, b = bh / AH, h = bh % AH, k0 = blockIdx.x * 32;
  const float* src = qkv + ((size_t)b * N + k0) * (3 * AE) + 2 * AE + h * AD;
  for (int e = threadIdx.x; e < 32 * 64; e += 256) t[e >> 6][e & 63] = src[(size_t)(e >> 6) * (3 * AE) + (e & 63)];   // rows of 64
  __syncthreads();
  for (int e = threadIdx.x; e < 64 * 32; e += 256) {
    const int n = e >> 5, k = e & 31;
    const float x = t[k][n];
    const float hv = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
    const size_t o = ((size_t)bh * 64 + n) * N + k0 + k;
    hi[o] = hv;
    lo[o] = (fabsf(x) <= 3.0e38f) ? x - hv : 0.f;        // inf / nan live in hi only
  }
  if (late && blockIdx.x == 0 && blockIdx.y == 0) asm volatile("griddepcontrol.wait;" ::: "memory");
}

struct PvArgs { float* out; int N, tiles_m, batch_heads; float* plane; int wide; };   // plane: mixed bf16 operand plane of out (or NULL); wide: out / plane rows are 32-byte aligned
__device__ __forceinline__ void st_global_v8(float* p, const uint32_t* v) {            // one 256-bit store (STG.E.256)
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]),
               "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}

// one head's 64 finished columns of this lane's row: result (merged heads) and, on request, the consumer's operand plane.
// A store instruction costs one L1 tag per row it touches (lane = row: 32 of them), so the wide form uses 256-bit stores (STG.256):
// half the instructions for the result, a quarter for the plane.  In attn_pv6_kernel this epilogue runs once per work item with
// nothing to overlap it: GPT-2 forward 8.82 -> 8.52 ms (profiles/r03X2_ab_gpt2_bench.log).
__device__ __forceinline__ void pv_store_head(const PvArgs& a, int b, int row, int h, const uint32_t (&v0)[32], const uint32_t (&v1)[32]) {
  float* crow = a.out + ((size_t)b * a.N + row) * AE + h * AD;
  if (a.wide) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      st_global_v8(crow + 8 * j, &v0[8 * j]);
      st_global_v8(crow + 32 + 8 * j, &v1[8 * j]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j)
      reinterpret_cast<float4*>(crow)[j] = make_float4(__uint_as_float(v0[4 * j]), __uint_as_float(v0[4 * j + 1]), __uint_as_float(v0[4 * j + 2]), __uint_as_float(v0[4 * j + 3]));
#pragma unroll
    for (int j = 0; j < 8; ++j)
      reinterpret_cast<float4*>(crow)[8 + j] = make_float4(__uint_as_float(v1[4 * j]), __uint_as_float(v1[4 * j + 1]), __uint_as_float(v1[4 * j + 2]), __uint_as_float(v1[4 * j + 3]));
  }
  if (!a.plane) return;                 // the projection that follows reads this row as its K-major A operand
  float* prow = a.plane + ((size_t)b * a.N + row) * AE;
  if (a.wide) {
    // 32 columns = two k-blocks of 16 = 128 contiguous bytes of plane: [16 bf16(x)][16 bf16(x - trunc x)] twice
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const uint32_t (&v)[32] = half ? v1 : v0;
      uint32_t w[32];
#pragma unroll
      for (int c = 0; c < 32; c += 2) {
        uint32_t h0, l0, h1, l1;
        plane_bits(__uint_as_float(v[c]), h0, l0);
        plane_bits(__uint_as_float(v[c + 1]), h1, l1);
        w[(c >> 4) * 16 + ((c & 15) >> 1)] = h0 | (h1 << 16);
        w[(c >> 4) * 16 + 8 + ((c & 15) >> 1)] = l0 | (l1 << 16);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) st_global_v8(prow + h * AD + half * 32 + 8 * j, &w[8 * j]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      plane_store4(prow, h * AD + 4 * j, __uint_as_float(v0[4 * j]), __uint_as_float(v0[4 * j + 1]), __uint_as_float(v0[4 * j + 2]), __uint_as_float(v0[4 * j + 3]));
      plane_store4(prow, h * AD + 32 + 4 * j, __uint_as_float(v1[4 * j]), __uint_as_float(v1[4 * j + 1]), __uint_as_float(v1[4 * j + 2]), __uint_as_float(v1[4 * j + 3]));
    }
  }
}

template <bool NORM>
__global__ void __launch_bounds__(P_THREADS, 1)
attn_pv_kernel(const __grid_constant__ CUtensorMap map_p, const __grid_constant__ CUtensorMap map_vh,
               const __grid_constant__ CUtensorMap map_vl, const __grid_constant__ CUtensorMap map_r, const __grid_constant__ PvArgs a) {
  constexpr int P_NS = p_ns(NORM), P_STAGE = p_stage(NORM);          // shadow the two-launch path's constants
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-byte aligned, and still known to be SHARED (LDS / STS, not generic LD / ST)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + P_NS * P_STAGE);
  uint64_t* full = bars;                   // [P_NS] TMA -> converters / MMA
  uint64_t* conv = bars + P_NS;            // [P_NS] converters -> MMA (4 arrivals, one per warp)
  uint64_t* empty = bars + 2 * P_NS;       // [P_NS] MMA -> TMA
  uint64_t* acc_full = bars + 3 * P_NS;    // [2]
  uint64_t* acc_empty = acc_full + 2;      // [2] (4 arrivals)
  uint64_t* afree = acc_empty + 2;         // [2] MMA -> converters: the P operand buffer in TMEM may be overwritten
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(afree + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = a.tiles_m * a.batch_heads;
  const int nkb = a.N / P_K;
  pdl_prologue();

  if (threadIdx.x == 0) {
    for (int s = 0; s < P_NS; ++s) { mbar_init(&full[s], 1); mbar_init(&conv[s], 4); mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 4); mbar_init(&afree[s], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      int s = 0; uint32_t ph = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int bh = t / a.tiles_m, mt = t % a.tiles_m;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* st = smem + s * P_STAGE;
          mbar_expect_tx(&full[s], P_STAGE);
          tma_load_3d(&map_p, &full[s], st, kb * P_K, mt * P_M, bh);
          tma_load_3d(&map_vh, &full[s], st + P_ATILE, kb * P_K, 0, bh);
          tma_load_3d(&map_vl, &full[s], st + P_ATILE + P_BTILE, kb * P_K, 0, bh);
          if (NORM) tma_load_3d(&map_r, &full[s], st + P_ATILE + 2 * P_BTILE, kb * P_K, mt * P_M, bh / AH);
          if (++s == P_NS) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      int s = 0; uint32_t ph = 0;
      int as = 0; uint32_t aph = 0, n = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        mbar_wait(&acc_empty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)(as * P_N);
        for (int kb = 0; kb < nkb; ++kb, ++n) {
          mbar_wait(&full[s], ph);          // the V planes (async proxy) ...
          mbar_wait(&conv[s], ph);          // ... and P's hi / lo planes in TMEM (tcgen05.st) are in place
          tc_fence_after();
          const uint32_t st = smem_u32(smem + s * P_STAGE);
          const uint64_t dbh = make_desc_sw128(st + P_ATILE), dbl = make_desc_sw128(st + P_ATILE + P_BTILE);
          const uint32_t ah = tmem_base + (uint32_t)(P_ACOL + (n & 1u) * 64u), al = ah + 32u;
#pragma unroll
          for (int k = 0; k < P_K / UK; ++k) {
            const uint64_t adv = (uint64_t)((k * UK * 4) >> 4);
            tc_mma_tf32_ts(d, al + k * UK, dbh + adv, P_IDESC, (kb | k) != 0);   // small terms first, then the main product
            tc_mma_tf32_ts(d, ah + k * UK, dbl + adv, P_IDESC, 1);
            tc_mma_tf32_ts(d, ah + k * UK, dbh + adv, P_IDESC, 1);
          }
          tc_commit(&empty[s]);
          tc_commit(&afree[n & 1u]);
          if (++s == P_NS) { s = 0; ph ^= 1; }
        }
        tc_commit(&acc_full[as]);
        if (++as == 2) { as = 0; aph ^= 1; }
      }
    }
  } else if (warp < 6) {
    // ---- converters: the P tile goes to TENSOR MEMORY as the A operand (row = lane, k = column): raw bits = hi plane,
    // lo = x - trunc(x); the MMAs then fetch only the V planes from shared memory
    const int q = warp & 3, row = q * 32 + lane;
    int s = 0; uint32_t ph = 0, n = 0, apar = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      for (int kb = 0; kb < nkb; ++kb, ++n) {
        const uint32_t u = n & 1u;
        mbar_wait(&full[s], ph);
        const uint8_t* st = smem + s * P_STAGE;
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) {                                // this lane's row: 8 chunks of 16 bytes, SWIZZLE_128B
          float4 x = *reinterpret_cast<const float4*>(st + row * 128 + ((j ^ (row & 7)) << 4));
          if (NORM) {      // p = RN(e r): the product the two-launch path forms in the score kernel's second pass
            const float4 r = *reinterpret_cast<const float4*>(st + P_ATILE + 2 * P_BTILE + row * 128 + ((j ^ (row & 7)) << 4));
            x.x = __fmul_rn(x.x, r.x); x.y = __fmul_rn(x.y, r.y); x.z = __fmul_rn(x.z, r.z); x.w = __fmul_rn(x.w, r.w);
          }
          hi[4 * j] = __float_as_uint(x.x); hi[4 * j + 1] = __float_as_uint(x.y); hi[4 * j + 2] = __float_as_uint(x.z); hi[4 * j + 3] = __float_as_uint(x.w);
          lo[4 * j] = __float_as_uint(tf32_lo(x.x)); lo[4 * j + 1] = __float_as_uint(tf32_lo(x.y));
          lo[4 * j + 2] = __float_as_uint(tf32_lo(x.z)); lo[4 * j + 3] = __float_as_uint(tf32_lo(x.w));
        }
        mbar_wait(&afree[u], ((apar >> u) & 1u) ^ 1u);               // the MMAs that read this buffer two k-blocks ago have retired
        apar ^= 1u << u;
        tc_fence_after();
        const uint32_t aa = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(P_ACOL + u * 64u);
        tc_st32(aa, hi);
        tc_st32(aa + 32, lo);
        tc_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&conv[s]);
        if (++s == P_NS) { s = 0; ph ^= 1; }
      }
    }
  } else {
    // ---- epilogue: accumulator -> out[b, row, h*64 .. h*64+63] (merged heads)
    const int q = warp & 3;
    int as = 0; uint32_t aph = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const int bh = t / a.tiles_m, mt = t % a.tiles_m, b = bh / AH, h = bh % AH;
      mbar_wait(&acc_full[as], aph);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * P_N);
      uint32_t v0[32], v1[32];
      tc_ld32(tbase, v0);
      tc_ld32(tbase + 32, v1);
      tc_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[as]);          // the accumulator is in registers now
      const int row = mt * P_M + q * 32 + lane;
      if (row < a.N) pv_store_head(a, b, row, h, v0, v1);
      if (++as == 2) { as = 0; aph ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(256) : "memory");
  }
}

// ------------------------------------------------------------------ P . V, six heads per CTA (wf2cu_attn_fused)
// attn_pv_kernel<true> works on one head at a time, so the reciprocal-denominator tile R[b, i-tile, j-tile] — shared by the 12 heads —
// is loaded again for every head: as many bytes through the TMA and the shared-memory pipe as E itself.  Here a CTA owns
// (batch, 128 queries, SIX heads): k-blocks outside, heads inside, six 64-column accumulators in tensor memory (384 columns + the two
// 64-column P operand buffers = 512), R loaded once per k-block into its own 2-slot ring and held in the converters' registers across the
// six heads.  Per head the MMAs run in the same order as attn_pv_kernel (k-blocks ascending; lo.hi, hi.lo, hi.hi), so the result is
// bit-identical to it.  B x tiles x 2 work items (128 at B = 8, N = 1024: one per CTA, the same 192 k-block steps as the busiest CTA before).
constexpr int P6_G = 6, P6_NS = 6, P6_ACOL = P6_G * P_N;                 // TMEM: [0, 384) accumulators, [384, 512) P operand buffers
constexpr int P6_RING = P6_NS * P_STAGE;                                 // E | V_hi | V_lo stages, then two R tiles
constexpr int P6_SMEM = P6_RING + 2 * P_ATILE + 1024 + 256;

__global__ void __launch_bounds__(P_THREADS, 1)
attn_pv6_kernel(const __grid_constant__ CUtensorMap map_p, const __grid_constant__ CUtensorMap map_vh,
                const __grid_constant__ CUtensorMap map_vl, const __grid_constant__ CUtensorMap map_r, const __grid_constant__ PvArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* rbuf = smem + P6_RING;
  uint64_t* bars = reinterpret_cast<uint64_t*>(rbuf + 2 * P_ATILE);
  uint64_t* full = bars;                   // [P6_NS] TMA -> converters / MMA
  uint64_t* conv = bars + P6_NS;           // [P6_NS] converters -> MMA (4 arrivals)
  uint64_t* empty = bars + 2 * P6_NS;      // [P6_NS] MMA -> TMA
  uint64_t* rfull = bars + 3 * P6_NS;      // [2] TMA -> converters
  uint64_t* rempty = rfull + 2;            // [2] converters -> TMA (4 arrivals)
  uint64_t* acc_full = rempty + 2;         // [1] MMA -> epilogue: all six accumulators of the item
  uint64_t* acc_empty = acc_full + 1;      // [1] epilogue -> MMA (4 arrivals)
  uint64_t* afree = acc_empty + 1;         // [2] MMA -> converters
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(afree + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = a.batch_heads / AH * a.tiles_m * (AH / P6_G);        // item = (b * tiles_m + mt) * 2 + head group
  const int nkb = a.N / P_K;
  pdl_prologue();

  if (threadIdx.x == 0) {
    for (int s = 0; s < P6_NS; ++s) { mbar_init(&full[s], 1); mbar_init(&conv[s], 4); mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&rfull[s], 1); mbar_init(&rempty[s], 4); mbar_init(&afree[s], 1); }
    mbar_init(acc_full, 1); mbar_init(acc_empty, 4);
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
      int s = 0; uint32_t ph = 0, rn = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int grp = t % (AH / P6_G), bm = t / (AH / P6_G), mt = bm % a.tiles_m, b = bm / a.tiles_m;
        for (int kb = 0; kb < nkb; ++kb, ++rn) {
          const uint32_t rs = rn & 1u;
          mbar_wait(&rempty[rs], ((rn >> 1) & 1u) ^ 1u);
          mbar_expect_tx(&rfull[rs], P_ATILE);
          tma_load_3d(&map_r, &rfull[rs], rbuf + rs * P_ATILE, kb * P_K, mt * P_M, b);
          for (int hh = 0; hh < P6_G; ++hh) {
            const int bh = b * AH + grp * P6_G + hh;
            mbar_wait(&empty[s], ph ^ 1);
            uint8_t* st = smem + s * P_STAGE;
            mbar_expect_tx(&full[s], P_STAGE);
            tma_load_3d(&map_p, &full[s], st, kb * P_K, mt * P_M, bh);
            tma_load_3d(&map_vh, &full[s], st + P_ATILE, kb * P_K, 0, bh);
            tma_load_3d(&map_vl, &full[s], st + P_ATILE + P_BTILE, kb * P_K, 0, bh);
            if (++s == P6_NS) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      int s = 0; uint32_t ph = 0, aph = 0, n = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        mbar_wait(acc_empty, aph ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < nkb; ++kb) {
          for (int hh = 0; hh < P6_G; ++hh, ++n) {
            mbar_wait(&full[s], ph);
            mbar_wait(&conv[s], ph);
            tc_fence_after();
            const uint32_t st = smem_u32(smem + s * P_STAGE);
            const uint64_t dbh = make_desc_sw128(st + P_ATILE), dbl = make_desc_sw128(st + P_ATILE + P_BTILE);
            const uint32_t d = tmem_base + (uint32_t)(hh * P_N);
            const uint32_t ah = tmem_base + (uint32_t)(P6_ACOL + (n & 1u) * 64u), al = ah + 32u;
#pragma unroll
            for (int k = 0; k < P_K / UK; ++k) {
              const uint64_t adv = (uint64_t)((k * UK * 4) >> 4);
              tc_mma_tf32_ts(d, al + k * UK, dbh + adv, P_IDESC, (kb | k) != 0);
              tc_mma_tf32_ts(d, ah + k * UK, dbl + adv, P_IDESC, 1);
              tc_mma_tf32_ts(d, ah + k * UK, dbh + adv, P_IDESC, 1);
            }
            tc_commit(&empty[s]);
            tc_commit(&afree[n & 1u]);
            if (++s == P6_NS) { s = 0; ph ^= 1; }
          }
        }
        tc_commit(acc_full);
        aph ^= 1;
      }
    }
  } else if (warp < 6) {
    const int q = warp & 3, row = q * 32 + lane;
    int s = 0; uint32_t ph = 0, n = 0, apar = 0, rn = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      for (int kb = 0; kb < nkb; ++kb, ++rn) {
        const uint32_t rs = rn & 1u;
        mbar_wait(&rfull[rs], (rn >> 1) & 1u);
        float4 r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = *reinterpret_cast<const float4*>(rbuf + rs * P_ATILE + row * 128 + ((j ^ (row & 7)) << 4));
        __syncwarp();
        if (lane == 0) mbar_arrive(&rempty[rs]);                     // the tile lives in registers for the six heads
        for (int hh = 0; hh < P6_G; ++hh, ++n) {
          const uint32_t u = n & 1u;
          mbar_wait(&full[s], ph);
          const uint8_t* st = smem + s * P_STAGE;
          uint32_t hi[32], lo[32];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 x = *reinterpret_cast<const float4*>(st + row * 128 + ((j ^ (row & 7)) << 4));
            x.x = __fmul_rn(x.x, r[j].x); x.y = __fmul_rn(x.y, r[j].y); x.z = __fmul_rn(x.z, r[j].z); x.w = __fmul_rn(x.w, r[j].w);
            hi[4 * j] = __float_as_uint(x.x); hi[4 * j + 1] = __float_as_uint(x.y); hi[4 * j + 2] = __float_as_uint(x.z); hi[4 * j + 3] = __float_as_uint(x.w);
            lo[4 * j] = __float_as_uint(tf32_lo(x.x)); lo[4 * j + 1] = __float_as_uint(tf32_lo(x.y));
            lo[4 * j + 2] = __float_as_uint(tf32_lo(x.z)); lo[4 * j + 3] = __float_as_uint(tf32_lo(x.w));
          }
          mbar_wait(&afree[u], ((apar >> u) & 1u) ^ 1u);
          apar ^= 1u << u;
          tc_fence_after();
          const uint32_t aa = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(P6_ACOL + u * 64u);
          tc_st32(aa, hi);
          tc_st32(aa + 32, lo);
          tc_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&conv[s]);
          if (++s == P6_NS) { s = 0; ph ^= 1; }
        }
      }
    }
  } else {
    const int q = warp & 3;
    uint32_t aph = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const int grp = t % (AH / P6_G), bm = t / (AH / P6_G), mt = bm % a.tiles_m, b = bm / a.tiles_m;
      mbar_wait(acc_full, aph);
      aph ^= 1;
      tc_fence_after();
      const int row = mt * P_M + q * 32 + lane;
#pragma unroll 1
      for (int hh = 0; hh < P6_G; ++hh) {
        const int h = grp * P6_G + hh;
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(hh * P_N);
        uint32_t v0[32], v1[32];
        tc_ld32(tbase, v0);
        tc_ld32(tbase + 32, v1);
        tc_ld_wait();
        if (row < a.N) pv_store_head(a, b, row, h, v0, v1);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

// ------------------------------------------------------------------ host side
static int encode(CUtensorMap* map, int rank, void* base, const cuuint64_t* dims, const cuuint64_t* strides, const cuuint32_t* box,
                  CUtensorMapSwizzle sw, CUtensorMapL2promotion l2, const char* what) {
  EncodeTiledFn enc = tensor_map_encoder();
  if (!enc) return WF2CU_ERR_CUDA;
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, l2,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(WF2CU_ERR_CUDA, "cuTensorMapEncodeTiled(%s) failed (%d)", what, (int)r);
  return 0;
}

static bool attn_shape_ok(int64_t batch, int64_t seq, int64_t n_head, int64_t head_dim) {
  return n_head == AH && head_dim == AD && batch >= 1 && batch <= 4096 && seq >= 128 && seq % 128 == 0 && seq <= (1 << 20);
}

}  // namespace wf2

using namespace wf2;

extern "C" {

int wf2cu_attn_heads_softmax(const void* qkv, int64_t qkv_off, int64_t batch, int64_t seq, int64_t n_head, int64_t head_dim,
                             float divisor, void* p, int64_t p_off) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(qkv && p, "wf2cu_attn_heads_softmax: null pointer");
  if (!attn_shape_ok(batch, seq, n_head, head_dim) || !(divisor > 0.f))
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_attn_heads_softmax: implemented for 12 heads of 64 features and seq a multiple of 128 "
                     "(got heads=%lld dim=%lld seq=%lld); issue the nodes separately", (long long)n_head, (long long)head_dim, (long long)seq);
  float* x = (float*)qkv + qkv_off;
  float* P = (float*)p + p_off;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(P) & 15))
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_attn_heads_softmax: operands must be 16-byte aligned");
  const int rc = attn_scores_raw(x, batch, seq, divisor, P);
  return rc;
}

// P . V for both entry points.  norm_r != NULL: P holds un-normalised exponentials and norm_r the (B N N) reciprocal denominators.
static int launch_pv(float* P, const float* norm_r, const float* x, float* vh, float* vl, int64_t batch, int64_t seq, float* O, const char* who) {
  const int N = (int)seq, BH = (int)(batch * AH);
  static const bool overlap = []{ const char* e = getenv("WF2CU_ATTN_SPLIT_OVERLAP"); return !e || atoi(e) != 0; }();   // 0: A/B
  const int late = (norm_r && overlap && pdl_enabled()) ? 1 : 0;      // only behind the score kernel of wf2cu_attn_fused
  WF2_CUDA(launch_pdl(attn_split_v_kernel, dim3((unsigned)(N / 32), (unsigned)BH), dim3(256), 0, S().stream, x, vh, vl, N, late));
  int rc = post_launch("wf2cu_attn_pv(split V)");
  if (rc) return rc;
  CUtensorMap mp, mvh, mvl, mr;
  {
    const cuuint64_t pd[3] = {(cuuint64_t)N, (cuuint64_t)N, (cuuint64_t)BH};
    const cuuint64_t ps[2] = {(cuuint64_t)N * 4, (cuuint64_t)N * N * 4};
    const cuuint32_t pb[3] = {P_K, P_M, 1};
    rc = encode(&mp, 3, P, pd, ps, pb, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "P");
    if (rc) return rc;
    const cuuint64_t vd[3] = {(cuuint64_t)N, (cuuint64_t)AD, (cuuint64_t)BH};
    const cuuint64_t vs[2] = {(cuuint64_t)N * 4, (cuuint64_t)AD * N * 4};
    const cuuint32_t vb[3] = {P_K, P_N, 1};
    rc = encode(&mvh, 3, vh, vd, vs, vb, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "V hi");
    if (rc) return rc;
    rc = encode(&mvl, 3, vl, vd, vs, vb, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "V lo");
    if (rc) return rc;
    mr = mp;
    if (norm_r) {
      const cuuint64_t rd[3] = {(cuuint64_t)N, (cuuint64_t)N, (cuuint64_t)batch};
      rc = encode(&mr, 3, (void*)norm_r, rd, ps, pb, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, "R");
      if (rc) return rc;
    }
  }
  static bool attr_set = false;
  if (!attr_set) {
    WF2_CUDA(cudaFuncSetAttribute(attn_pv_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, p_smem(false)));
    WF2_CUDA(cudaFuncSetAttribute(attn_pv_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, p_smem(true)));
    WF2_CUDA(cudaFuncSetAttribute(attn_pv6_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, P6_SMEM));
    attr_set = true;
  }
  PvArgs args{O, N, N / P_M, BH, nullptr, 0};
  if (S().emit_next) args.plane = planes_emit_begin(batch * seq, AE);     // wf2cu_emit_operand_planes: one-shot, may decline
  args.wide = ((reinterpret_cast<uintptr_t>(O) & 31) == 0 && (reinterpret_cast<uintptr_t>(args.plane) & 31) == 0) ? 1 : 0;
  const int total = args.tiles_m * BH;
  const int grid = total < S().sm_count ? total : S().sm_count;
  static const bool six = []{ const char* e = getenv("WF2CU_ATTN_PV6"); return !e || atoi(e) != 0; }();   // 0: one head per work item (A/B)
  if (norm_r && six) {
    const int items = (int)batch * args.tiles_m * (AH / P6_G);
    WF2_CUDA(launch_pdl(attn_pv6_kernel, dim3(items < S().sm_count ? items : S().sm_count), dim3(P_THREADS), P6_SMEM, S().stream, mp, mvh, mvl, mr, args));
  } else if (norm_r) WF2_CUDA(launch_pdl(attn_pv_kernel<true>, dim3(grid), dim3(P_THREADS), p_smem(true), S().stream, mp, mvh, mvl, mr, args));
  else        WF2_CUDA(launch_pdl(attn_pv_kernel<false>, dim3(grid), dim3(P_THREADS), p_smem(false), S().stream, mp, mvh, mvl, mr, args));
  rc = post_launch(who);
  if (!rc && args.plane) planes_emit_publish(O, batch * seq, AE, AE, args.plane);
  return rc;
}

int wf2cu_attn_pv(const void* p, int64_t p_off, const void* qkv, int64_t qkv_off, int64_t batch, int64_t seq, int64_t n_head,
                  int64_t head_dim, void* out, int64_t out_off) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(qkv && p && out, "wf2cu_attn_pv: null pointer");
  if (!attn_shape_ok(batch, seq, n_head, head_dim) || S().branch != 0)
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_attn_pv: implemented for 12 heads of 64 features and seq a multiple of 128 on the main "
                     "stream (got heads=%lld dim=%lld seq=%lld); issue the nodes separately", (long long)n_head, (long long)head_dim, (long long)seq);
  float* P = (float*)p + p_off;
  const float* x = (const float*)qkv + qkv_off;
  float* O = (float*)out + out_off;
  if ((reinterpret_cast<uintptr_t>(P) & 15) || (reinterpret_cast<uintptr_t>(O) & 15))
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_attn_pv: operands must be 16-byte aligned");
  const size_t plane = (size_t)(batch * AH) * AD * seq * 4;
  float* vh = tc_workspace(2 * plane);
  if (!vh) return WF2CU_ERR_CUDA;
  float* vl = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(vh) + plane);
  return launch_pv(P, nullptr, x, vh, vl, batch, seq, O, "wf2cu_attn_pv");
}

int wf2cu_attn_fused(const void* qkv, int64_t qkv_off, int64_t batch, int64_t seq, int64_t n_head, int64_t head_dim, float divisor,
                     void* out, int64_t out_off) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(qkv && out, "wf2cu_attn_fused: null pointer");
  if (!attn_shape_ok(batch, seq, n_head, head_dim) || !(divisor > 0.f) || S().branch != 0)
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_attn_fused: implemented for 12 heads of 64 features and seq a multiple of 128 on the main "
                     "stream (got heads=%lld dim=%lld seq=%lld); issue the nodes separately", (long long)n_head, (long long)head_dim, (long long)seq);
  const float* x = (const float*)qkv + qkv_off;
  float* O = (float*)out + out_off;
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(O) & 15))
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_attn_fused: operands must be 16-byte aligned");
  // workspace: E (B H N N) | R (B N N) | V_hi | V_lo
  const size_t ebytes = (size_t)batch * AH * seq * seq * 4, rbytes = (size_t)batch * seq * seq * 4, plane = (size_t)(batch * AH) * AD * seq * 4;
  float* ws = tc_workspace(ebytes + rbytes + 2 * plane);
  if (!ws) return WF2CU_ERR_CUDA;
  float* E = ws;
  float* R = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ws) + ebytes);
  float* vh = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(R) + rbytes);
  float* vl = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(vh) + plane);
  const int emit = S().emit_next;            // the request is for the P.V launch: keep it across the score kernel
  S().emit_next = 0;
  int rc = attn_scores_onepass(x, batch, seq, divisor, E, R);
  if (rc) return rc;
  S().emit_next = emit;
  return launch_pv(E, R, x, vh, vl, batch, seq, O, "wf2cu_attn_fused");
}

}  // extern "C"
