// gemm_raw.cuh — the tcgen05 3xTF32 mainloop WITHOUT an operand pre-pass (included by gemm_tc.cu after its epilogue helpers).
//
// gemm_tc_kernel / gemm_tc256_kernel read hi / lo planes that split_planes_* wrote to global memory first: two extra
// launches, 2 x (read 4 + write 8) bytes per operand element and twice the L2 -> SM bytes in the main loop (ncu, round 1:
// 744 MB of DRAM traffic against 201 MB algorithmic at N = 4096).  Here a K-major operand is loaded by TMA as RAW fp32:
//   * kind::tf32 reads the upper 19 bits of a 32-bit container, so the raw tile IS the hi plane (truncation, checked on
//     hardware by scripts/micro/ts_mma_test.cu and by the error of every test: ~1e-6);
//   * lo = x - trunc(x) is element-wise and the swizzled shared-memory layout is position preserving, so four converter
//     warps write the lo tile at the same offsets of a second buffer (generic-proxy stores + fence.proxy.async).
// An operand that is MN-major for the product (B of a plain `(!matmul a b)`: (k n) row-major) still goes through the
// transposing split (it has to be transposed anyway); its planes arrive as before (B_RAW = false).
//
// One persistent CTA per SM, 128 x 256 tiles, k-blocks of 16 (64-byte rows, SWIZZLE_64B), 3 stages of 48 KiB,
// two 256-column accumulators in TMEM so the epilogue of a tile overlaps the main loop of the next:
//   warp 0      TMA producer       warp 1      MMA issuer (6 x tcgen05.mma 128x256x8 per k-block)
//   warps 2-5   converters         warps 6-..  epilogue (4 x CG warps: TMEM lane quarter x column group)
// Measured on B200 (scripts/micro/mma_rate_test.cu; see the correction in DESIGN.md 3.5: the 45 was the issue loop of `lane == 0`,
// with A in shared memory the floor is ~40 cycles, with A in tensor memory N / 2 all the way down): a kind::tf32 MMA with M = 128 costs max(45, N / 2) cycles — N = 256 runs
// at the nominal rate (128 cycles), anything below N = 96 pays the 45-cycle floor — so tiles are as wide as the instruction
// allows.  That is also why the GPT-2 example's attention scores run through this main loop (N = 256 keys per MMA) instead of
// keeping all 12 heads of a narrow tile in TMEM (N = 32 per MMA, 288 MMAs per 128 x 32 tile: 271 us per layer).
//
// The work list and the epilogue are template parameters:
//   RawGemmJob      MatmulNode: C = op(A) . op(B) (+ bias, relu), C through bulk tensor stores
//   RawAttnJob      the GPT-2 example's attention weights: per (batch, 128 queries, 256 keys) 24 sub-tiles — the 12 heads once
//                   to accumulate den = sum_h exp((Q_h . K_h) / N) in registers (head order), then the 12 heads again to write
//                   P = exp(...) / den.  Raw operands + converter warps by default (194 us per layer at B = 8, N = 1024);
//                   WF2CU_ATTN_PLANES=1 pre-splits the lo planes instead (184 us + a 16 us split launch).  Ablation
//                   (WF2CU_ATTN_DBG): main loop alone 147 us, pass-2 arithmetic + staging +26, P stores +22.
#pragma once

namespace wf2 {

constexpr int RM = 128, RN = 256, RK = 16;
constexpr int R_ATILE = RM * RK * 4;                 // 8 KiB
constexpr int R_BTILE = RN * RK * 4;                 // 16 KiB
constexpr int R_STAGE = 2 * R_ATILE + 2 * R_BTILE;   // A raw | A lo | B raw (or hi) | B lo = 48 KiB
constexpr int R_NS = 3;
constexpr int R_CONV_WARPS = 4;
constexpr uint32_t R_IDESC = idesc_tf32_mn(RM, RN), R_IDESC_B = idesc_bf16_mn(RM, RN);
// converter warps exist only when an operand arrives raw: idle warps would still take their share of the register file
__host__ __device__ constexpr int raw_threads(int epi_warps, bool convert) { return 64 + (convert ? 32 * R_CONV_WARPS : 0) + 32 * epi_warps; }
__host__ __device__ constexpr int raw_smem(int epi_warps, int epi_bytes_per_warp) { return R_NS * R_STAGE + epi_warps * epi_bytes_per_warp + 1024 + 256; }

struct RawCoords {            // TMA coordinates of one sub-tile's operands and its k extent
  int a_k0, a_row, a_b;
  int b_k0, b_row, b_b;
  int nkb;                    // k-blocks of 16
};

__device__ __forceinline__ float raw_lo(float x) {   // x - trunc_tf32(x), exact; inf / nan live in the hi part only
  return (fabsf(x) <= 3.0e38f) ? x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u) : 0.f;
}

// the mixed bf16 lo tile of a raw operand (same bytes as the fp32 lo tile): per 64-byte row bf16_rn(x[0:16]) then
// bf16_rn(x[0:16] - trunc_tf32(x[0:16])); inf / nan contribute 0 to the corrections
__device__ __forceinline__ uint32_t raw_bf2(float a, float b) {
  return (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(a)) | ((uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(b)) << 16);
}
__device__ __forceinline__ float raw_fin(float x) { return (fabsf(x) <= 3.0e38f) ? x : 0.f; }

// ---------------------------------------------------------------- MatmulNode
struct RawGemmJob {
  static constexpr int kEpiWarps = 4, kEpiBytes = EPI_SLOT, kSubTiles = 1;
  TcArgs t;                   // C, ldc, m, n, tiles, bias, act, c_tma ... (nkb in k-blocks of 16)
  __device__ __forceinline__ bool hyb() const { return t.passes == 2; }   // corrections as bf16 MMAs on mixed lo tiles
  __device__ __forceinline__ int items() const { return t.tiles_m * t.tiles_n * t.batch; }
  __device__ __forceinline__ RawCoords coords(int w, int) const {
    int b, mt, nt;
    tile_coords(w, t, b, mt, nt);
    return RawCoords{0, mt * RM, t.a_bcast ? 0 : b, 0, nt * RN, t.b_bcast ? 0 : b, t.nkb};
  }
};

// ---------------------------------------------------------------- attention scores (examples/gpt-2/utils.lisp:222-244)
constexpr int RA_H = 12, RA_D = 64, RA_E = RA_H * RA_D;
struct RawAttnArgs {
  int B, N, tiles_i, tiles_j;  // tiles of 128 queries x 256 keys
  float scale;                 // log2(e) / divisor
  int dbg;                     // measurement switches (WF2CU_ATTN_DBG): 1 no P stores, 2 no pass-2 arithmetic, 4 no pass-1 arithmetic
  int hyb;                     // corrections as bf16 MMAs on mixed lo tiles (see split_bf in gemm_tc.cu)
};
struct RawAttnJob {
  static constexpr int kEpiWarps = 16, kEpiBytes = 2 * 1024, kSubTiles = 2 * RA_H;
  RawAttnArgs a;
  __device__ __forceinline__ bool hyb() const { return a.hyb != 0; }
  __device__ __forceinline__ int items() const { return a.B * a.tiles_i * a.tiles_j; }
  __device__ __forceinline__ void where(int w, int& jt, int& it, int& b) const {
    jt = w % a.tiles_j; it = (w / a.tiles_j) % a.tiles_i; b = w / (a.tiles_j * a.tiles_i);
  }
  __device__ __forceinline__ RawCoords coords(int w, int sub) const {
    int jt, it, b;
    where(w, jt, it, b);
    const int h = sub < RA_H ? sub : sub - RA_H;
    return RawCoords{h * RA_D, it * RM, b, RA_E + h * RA_D, jt * RN, b, RA_D / RK};
  }
};

// One-pass form (wf2cu_attn_fused): the 12 heads ONCE — E[b,h,i,j] = exp(s_h / N) leaves un-normalised, den accumulates in registers as
// above, and after the twelfth head R[b,i,j] = RN(1 / den) goes out as a (B N N) tensor; the P.V kernel multiplies E by R on its way
// into tensor memory (p = RN(e r), the very product the second pass forms).  Half the Q.K^T work, one exp per element instead of two.
// R leaves through the tensor map in the `map_al` slot (unused when both operands arrive raw).
struct RawAttnJob1 {
  static constexpr int kEpiWarps = 16, kEpiBytes = 2 * 1024, kSubTiles = RA_H;
  RawAttnArgs a;
  __device__ __forceinline__ bool hyb() const { return false; }
  __device__ __forceinline__ int items() const { return a.B * a.tiles_i * a.tiles_j; }
  __device__ __forceinline__ void where(int w, int& jt, int& it, int& b) const {
    jt = w % a.tiles_j; it = (w / a.tiles_j) % a.tiles_i; b = w / (a.tiles_j * a.tiles_i);
  }
  __device__ __forceinline__ RawCoords coords(int w, int sub) const {
    int jt, it, b;
    where(w, jt, it, b);
    return RawCoords{sub * RA_D, it * RM, b, RA_E + sub * RA_D, jt * RN, b, RA_D / RK};
  }
};

// exp(s / divisor) = 2^(s * log2(e) / divisor): one multiply by the pre-rounded constant and ex2.approx (relative error <= 2^-22
// + 0.7 * 2^-23 |s / divisor|: ~3 ulp for the |s / divisor| < 1 the division by N produces).  The epilogue is what bounds this
// kernel: with the library's 1-ulp exp (10 instructions per element, twice per element because the product is recomputed) it
// ran in 178 us per layer; the weights feed a softmax whose parity bar is 1e-5 relative (BASELINE.json), 40 x above this.
template <bool POW2>
__device__ __forceinline__ float raw_attn_exp(float s, float scale) {
  float e;
  asm("ex2.approx.f32 %0, %1;" : "=f"(e) : "f"(__fmul_rn(s, scale)));
  return e;
}

// JOB: one of the structs above; A_RAW / B_RAW: the operand arrives as raw fp32 (K-major) and its lo tile is made here by the
// converter warps, else its lo plane comes from a second tensor map; POW2: see raw_attn_exp (attention job only)
template <class JOB, bool A_RAW, bool B_RAW, bool POW2>
__global__ void __launch_bounds__(raw_threads(JOB::kEpiWarps, A_RAW || B_RAW), 1)
gemm_raw_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_al,
                const __grid_constant__ CUtensorMap map_b, const __grid_constant__ CUtensorMap map_bl,
                const __grid_constant__ CUtensorMap map_c, const __grid_constant__ JOB job) {
  constexpr bool CONVERT = A_RAW || B_RAW;
  constexpr int NCONV = CONVERT ? R_CONV_WARPS : 0;
  constexpr int EW = JOB::kEpiWarps, CG = EW / 4, COLS = RN / CG;    // epilogue warps, column groups, columns per warp
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-byte aligned, and still known to be SHARED (LDS / STS, not generic LD / ST)
  uint8_t* epi_base = smem + R_NS * R_STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_base + EW * JOB::kEpiBytes);
  uint64_t* full = bars;                     // [R_NS] TMA -> converters, MMA
  uint64_t* empty = bars + R_NS;             // [R_NS] MMA -> TMA
  uint64_t* conv = bars + 2 * R_NS;          // [R_NS] converters -> MMA (one arrival per converter warp)
  uint64_t* acc_full = bars + 3 * R_NS;      // [2]    MMA -> epilogue
  uint64_t* acc_empty = acc_full + 2;        // [2]    epilogue -> MMA (EW arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int items = job.items();
  // bf16 corrections only on pre-split (mixed) lo planes: the converter warps are the raw variants' bottleneck, and mixed
  // tiles cost them two scattered 8-byte stores per chunk instead of one 16-byte store (r02f: 194 -> 336 us on attention)
  const bool hyb = (A_RAW || B_RAW) ? false : job.hyb();
  if constexpr (std::is_same<JOB, RawAttnJob1>::value) {
    // wait FIRST, then let the dependent grid in: the kernel launched after this one in wf2cu_attn_fused (the split of V) does not
    // depend on this kernel at all, only on qkv — it starts as soon as every CTA here has passed this point, i.e. once the producer
    // of qkv is known to be complete, and runs beside the score tiles (attn_split_v_kernel, `late`)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  } else {
    pdl_prologue();
  }

  if (threadIdx.x == 0) {
    for (int s = 0; s < R_NS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); mbar_init(&conv[s], CONVERT ? R_CONV_WARPS : 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], EW); }
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
    // ===================== TMA producer =====================
    if (elect_one()) {
      constexpr uint32_t tx_bytes = R_ATILE + R_BTILE + (A_RAW ? 0 : R_ATILE) + (B_RAW ? 0 : R_BTILE);
      int s = 0; uint32_t ph = 0;
      for (int w = blockIdx.x; w < items; w += gridDim.x) {
        for (int sub = 0; sub < JOB::kSubTiles; ++sub) {
          const RawCoords c = job.coords(w, sub);
          for (int kb = 0; kb < c.nkb; ++kb) {
            mbar_wait(&empty[s], ph ^ 1);
            uint8_t* st = smem + s * R_STAGE;
            mbar_expect_tx(&full[s], tx_bytes);
            tma_load_3d(&map_a, &full[s], st, c.a_k0 + kb * RK, c.a_row, c.a_b);
            if (!A_RAW) tma_load_3d(&map_al, &full[s], st + R_ATILE, c.a_k0 + kb * RK, c.a_row, c.a_b);
            tma_load_3d(&map_b, &full[s], st + 2 * R_ATILE, c.b_k0 + kb * RK, c.b_row, c.b_b);
            if (!B_RAW) tma_load_3d(&map_bl, &full[s], st + 2 * R_ATILE + R_BTILE, c.b_k0 + kb * RK, c.b_row, c.b_b);
            if (++s == R_NS) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (elect_one()) {
      int s = 0; uint32_t ph = 0;
      int as = 0; uint32_t aph = 0;
      for (int w = blockIdx.x; w < items; w += gridDim.x) {
        for (int sub = 0; sub < JOB::kSubTiles; ++sub) {
          const int nkb = job.coords(w, sub).nkb;
          mbar_wait(&acc_empty[as], aph ^ 1);     // the epilogue has drained this accumulator
          tc_fence_after();
          const uint32_t d = tmem_base + (uint32_t)(as * RN);
          for (int kb = 0; kb < nkb; ++kb) {
            mbar_wait(&full[s], ph);              // tiles loaded by TMA (async proxy) ...
            if (CONVERT) mbar_wait(&conv[s], ph); // ... and lo tiles made here (generic proxy + fence.proxy.async) are in place
            tc_fence_after();
            const uint32_t st = smem_u32(smem + s * R_STAGE);
            const uint64_t dah = make_desc_sw64(st), dal = make_desc_sw64(st + R_ATILE);
            const uint64_t dbh = make_desc_sw64(st + 2 * R_ATILE), dbl = make_desc_sw64(st + 2 * R_ATILE + R_BTILE);
            if (hyb) {
              // the 64-byte swizzle row of a mixed lo tile: hb[0:16] then lb[0:16]; bf16(A) . bf16(B_lo) + bf16(A_lo) . bf16(B)
              // as two bf16 MMAs of K = 16, then the main product on the raw tiles (the tensor core truncates fp32 to TF32)
              tc_mma_bf16(d, dal, dbl + 2, R_IDESC_B, kb != 0);
              tc_mma_bf16(d, dal + 2, dbl, R_IDESC_B, 1);
              tc_mma_tf32(d, dah, dbh, R_IDESC, 1);
              tc_mma_tf32(d, dah + 2, dbh + 2, R_IDESC, 1);
            } else {
#pragma unroll
              for (int k = 0; k < RK / UK; ++k) {
                const uint64_t adv = (uint64_t)((k * UK * 4) >> 4);
                tc_mma_tf32(d, dal + adv, dbh + adv, R_IDESC, (kb | k) != 0);   // small terms first, then the main product
                tc_mma_tf32(d, dah + adv, dbl + adv, R_IDESC, 1);
                tc_mma_tf32(d, dah + adv, dbh + adv, R_IDESC, 1);
              }
            }
            tc_commit(&empty[s]);
            if (++s == R_NS) { s = 0; ph ^= 1; }
          }
          tc_commit(&acc_full[as]);
          if (++as == 2) { as = 0; aph ^= 1; }
        }
      }
    }
  } else if (warp < 2 + NCONV) {
    // ===================== converters: lo = x - trunc_tf32(x), same offset in the lo tile =====================
    const int ct = threadIdx.x - 64;                                 // 0..127
    constexpr int CHUNKS = ((A_RAW ? R_ATILE : 0) + (B_RAW ? R_BTILE : 0)) / 16;   // 512 of A, 1024 of B
    constexpr int FIRST = A_RAW ? 0 : R_ATILE / 16;                  // chunk numbering: [0, 512) A, [512, 1536) B
    int s = 0; uint32_t ph = 0;
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
      for (int sub = 0; sub < JOB::kSubTiles; ++sub) {
        const int nkb = job.coords(w, sub).nkb;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full[s], ph);
          uint8_t* st = smem + s * R_STAGE;
#pragma unroll
          for (int i = 0; i < CHUNKS / 128; ++i) {
            const int c = FIRST + ct + i * 128;
            const int src = c < R_ATILE / 16 ? c * 16 : 2 * R_ATILE + (c - R_ATILE / 16) * 16;
            const int dst = src + (c < R_ATILE / 16 ? R_ATILE : R_BTILE);
            const float4 x = *reinterpret_cast<const float4*>(st + src);
            *reinterpret_cast<float4*>(st + dst) = make_float4(raw_lo(x.x), raw_lo(x.y), raw_lo(x.z), raw_lo(x.w));
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(&conv[s]);
          if (++s == R_NS) { s = 0; ph ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue =====================
    const int ew = warp - 2 - NCONV, q = warp & 3, cg = ew >> 2;   // TMEM lane quarter (= warp % 4, fixed by the hardware), column group
    const bool leader = elect_one();                               // the lane that issues (and waits for) this warp's bulk stores
    uint8_t* slot = epi_base + ew * JOB::kEpiBytes;
    uint32_t which = 0;
    int as = 0; uint32_t aph = 0;
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
      if constexpr (std::is_same<JOB, RawGemmJob>::value) {
        // ---- MatmulNode: TMEM -> registers -> (bias, relu) -> staging block -> bulk tensor store, as gemm_tc_kernel
        const TcArgs& a = job.t;
        int b, mt, nt;
        tile_coords(w, a, b, mt, nt);
        mbar_wait(&acc_full[as], aph);
        tc_fence_after();
        float* cbase = a.C + (int64_t)b * a.sc;
        float* stage = reinterpret_cast<float*>(slot);
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * RN);
        uint32_t v0[32], v1[32];
        tc_ld32(tbase, v0);
#pragma unroll 1
        for (int c = 0; c < RN / 32; c += 2) {
          tc_ld_wait();
          tc_ld32(tbase + (c + 1) * 32, v1);
          int col0 = nt * RN + c * 32;
          if (col0 < a.n) tc_epilogue32(a, v0, mt * RM + q * 32, col0, lane);
          if (a.c_tma) tc_store32_tma(a, &map_c, slot, which, v0, mt * RM + q * 32, col0, b, lane);
          else tc_store32(a, cbase, stage, v0, mt * RM + q * 32, col0, lane);
          tc_ld_wait();
          if (c + 2 < RN / 32) tc_ld32(tbase + (c + 2) * 32, v0);
          col0 += 32;
          if (col0 < a.n) tc_epilogue32(a, v1, mt * RM + q * 32, col0, lane);
          if (a.c_tma) tc_store32_tma(a, &map_c, slot, which, v1, mt * RM + q * 32, col0, b, lane);
          else tc_store32(a, cbase, stage, v1, mt * RM + q * 32, col0, lane);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_empty[as]);
        if (++as == 2) { as = 0; aph ^= 1; }
      } else if constexpr (std::is_same<JOB, RawAttnJob1>::value) {
        // ---- one pass: E = exp(s_h / N) out per head, den in registers, R = RN(1 / den) out after the last head
        const RawAttnArgs& a = job.a;
        int jt, it, b;
        job.where(w, jt, it, b);
        const int col0 = jt * RN + cg * COLS;
        float den[COLS];
#pragma unroll
        for (int c = 0; c < COLS; ++c) den[c] = 0.f;
#pragma unroll 1
        for (int h = 0; h < RA_H; ++h) {
          mbar_wait(&acc_full[as], aph);
          tc_fence_after();
          const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * RN + cg * COLS);
#pragma unroll
          for (int c0 = 0; c0 < COLS; c0 += 8) {
            uint32_t v[8];
            tc_ld8(tbase + c0, v);
            uint8_t* blk = slot + which * 1024;
            which ^= 1;
            if (leader) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            tc_ld_wait();
            if (c0 + 8 >= COLS) {                                   // the accumulator is in registers now
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&acc_empty[as]);
            } else {
              __syncwarp();
            }
            float e[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              e[c] = raw_attn_exp<POW2>(__uint_as_float(v[c]), a.scale);
              den[c0 + c] = __fadd_rn(den[c0 + c], e[c]);
            }
            uint8_t* mine = blk + lane * 32;                       // SWIZZLE_32B: 16-byte chunk j of row r at j ^ ((r >> 2) & 1)
            *reinterpret_cast<float4*>(mine + ((0 ^ ((lane >> 2) & 1)) << 4)) = make_float4(e[0], e[1], e[2], e[3]);
            *reinterpret_cast<float4*>(mine + ((1 ^ ((lane >> 2) & 1)) << 4)) = make_float4(e[4], e[5], e[6], e[7]);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (leader) tma_store_4d(&map_c, blk, col0 + c0, it * RM + q * 32, h, b);
          }
          if (++as == 2) { as = 0; aph ^= 1; }
        }
#pragma unroll
        for (int c0 = 0; c0 < COLS; c0 += 8) {
          uint8_t* blk = slot + which * 1024;
          which ^= 1;
          if (leader) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          uint8_t* mine = blk + lane * 32;
          *reinterpret_cast<float4*>(mine + ((0 ^ ((lane >> 2) & 1)) << 4)) =
              make_float4(__frcp_rn(den[c0]), __frcp_rn(den[c0 + 1]), __frcp_rn(den[c0 + 2]), __frcp_rn(den[c0 + 3]));
          *reinterpret_cast<float4*>(mine + ((1 ^ ((lane >> 2) & 1)) << 4)) =
              make_float4(__frcp_rn(den[c0 + 4]), __frcp_rn(den[c0 + 5]), __frcp_rn(den[c0 + 6]), __frcp_rn(den[c0 + 7]));
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (leader) tma_store_3d(&map_al, blk, col0 + c0, it * RM + q * 32, b);
        }
      } else {
        // ---- the GPT-2 example's attention weights: 12 head sub-tiles to build den = sum_h exp(s_h / N) in registers (head
        // order, like `!sum`), then the 12 heads again: p = exp(s_h / N) / den, 32 x 8 blocks through the bulk-copy engine
        const RawAttnArgs& a = job.a;
        int jt, it, b;
        job.where(w, jt, it, b);
        const int col0 = jt * RN + cg * COLS;
        float den[COLS];
#pragma unroll
        for (int c = 0; c < COLS; ++c) den[c] = 0.f;
#pragma unroll 1
        for (int sub = 0; sub < 2 * RA_H; ++sub) {
          mbar_wait(&acc_full[as], aph);
          tc_fence_after();
          const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * RN + cg * COLS);
          if (sub < RA_H) {
#pragma unroll
            for (int c0 = 0; c0 < COLS; c0 += 16) {
              uint32_t v[16];
              tc_ld16(tbase + c0, v);
              tc_ld_wait();
#pragma unroll
              if (!(a.dbg & 4))
#pragma unroll
                for (int c = 0; c < 16; ++c) den[c0 + c] = __fadd_rn(den[c0 + c], raw_attn_exp<POW2>(__uint_as_float(v[c]), a.scale));
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[as]);
            if (sub == RA_H - 1) {
              // from here on den[] holds r = RN(1 / den): p = RN(e r) is within 1 ulp of the IEEE quotient e / den
#pragma unroll
              for (int c = 0; c < COLS; ++c) den[c] = __frcp_rn(den[c]);
            }
          } else {
            const int h = sub - RA_H;
#pragma unroll
            for (int c0 = 0; c0 < COLS; c0 += 8) {
              uint32_t v[8];
              tc_ld8(tbase + c0, v);
              uint8_t* blk = slot + which * 1024;
              which ^= 1;
              if (leader) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
              tc_ld_wait();
              if (c0 + 8 >= COLS) {                                   // the accumulator is in registers now
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[as]);
              } else {
                __syncwarp();
              }
              if (a.dbg & 2) continue;
              float p[8];
#pragma unroll
              for (int c = 0; c < 8; ++c) p[c] = __fmul_rn(raw_attn_exp<POW2>(__uint_as_float(v[c]), a.scale), den[c0 + c]);
              uint8_t* mine = blk + lane * 32;                       // SWIZZLE_32B: 16-byte chunk j of row r at j ^ ((r >> 2) & 1)
              *reinterpret_cast<float4*>(mine + ((0 ^ ((lane >> 2) & 1)) << 4)) = make_float4(p[0], p[1], p[2], p[3]);
              *reinterpret_cast<float4*>(mine + ((1 ^ ((lane >> 2) & 1)) << 4)) = make_float4(p[4], p[5], p[6], p[7]);
              asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
              __syncwarp();
              if (leader && !(a.dbg & 1)) tma_store_4d(&map_c, blk, col0 + c0, it * RM + q * 32, h, b);
            }
          }
          if (++as == 2) { as = 0; aph ^= 1; }
        }
      }
    }
    if (JOB::kEpiBytes > 0 && (leader || lane == 0)) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

}  // namespace wf2
