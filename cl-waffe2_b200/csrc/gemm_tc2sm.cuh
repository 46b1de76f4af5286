// gemm_tc2sm.cuh — the plane-fed tcgen05 kernel on a CTA PAIR (cta_group::2); included by gemm_tc.cu after gemm_tc256_kernel.
//
// Why: with the corrections as bf16 MMAs a k-block of 16 costs the tensor pipe 1024 cycles per 256 x 256 tile, but the tile needs
// 64 KiB of operand planes for it = 64 B/clk per SM, more than the L2 delivers to one SM (measured: 1470 cycles per k-block,
// 44.6 B/clk; the 128 x 256 and 256 x 192 shapes sit at 56 and 48 B/clk).  The single-CTA kernels are bound by OPERAND BYTES, not
// by the tensor pipe.  A CTA pair executes one MMA of M = 256 across two SMs (128 rows each) and BOTH read the B operand from
// the two shared memories: each CTA loads only HALF of the B tile.  Pair tile 512 x 256 (every CTA owns 256 rows: two M = 256 pair
// MMAs per k-step, accumulators 2 x 256 columns = all of TMEM as in gemm_tc256_kernel):
//     per CTA and k-block of 16:  A 256 x 64 B x 2 planes = 32 KiB  +  B 128 x 64 B x 2 planes = 16 KiB  = 48 KiB (was 64)
// Protocol (cluster of 2 along x; rank 0 = leader):
//   warp 0 of BOTH CTAs   TMA producer: waits for its own empty[s] (multicast commit of the leader), then cp.async.bulk.tensor
//                          ... .cta_group::2 whose bytes complete on the LEADER's full[s]; the leader alone arrives (expect_tx of
//                          both CTAs' bytes)
//   warp 1 of the leader  MMA issuer: tcgen05.mma.cta_group::2 (M = 256, N = 256), tcgen05.commit ... multicast::cluster to
//                          empty[s] / acc_full of both CTAs
//   warps 2-9 of both     epilogue of the CTA's own 256 rows, two warps per TMEM lane quarter (128 of the 256 columns each): the two
//                          accumulators fill tensor memory, so the epilogue is NOT overlapped with the next tile's main loop and at K = 768
//                          (GPT-2's projections) it was a third of the tile; acc_empty lives in the leader (16 arrivals, the peer's remote)
// Bounded waits (mbar_wait traps) turn a protocol bug into a CUDA error instead of a hung GPU.
#pragma once

namespace wf2 {

constexpr int P2_STAGES = 4;
constexpr int P2_APLANE = 256 * T2K * 4;            // 16 KiB: 256 rows x 64 B
constexpr int P2_BHALF = 128 * T2K * 4;             // 8 KiB: this CTA's half of the 256 B rows
constexpr int P2_STAGE = 2 * P2_APLANE + 2 * P2_BHALF;   // A_hi A_lo B_hi B_lo = 48 KiB
constexpr int P2_EPI_WARPS = EPI8_WARPS, P2_THREADS = EPI8_THREADS, P2_EPI_SLOT = EPI8_SLOT;
constexpr int P2_SMEM = P2_STAGES * P2_STAGE + 1024 + EPI_BYTES + 256;
static_assert(P2_SMEM <= 232448 && P2_STAGE % 1024 == 0, "shared-memory budget");
constexpr uint32_t kIdescTf32P2 = idesc_tf32_mn(256, 256), kIdescBf16P2 = idesc_bf16_mn(256, 256);

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
  uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx_cluster(uint32_t cluster_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.release.cluster.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_3d_2sm(const CUtensorMap* map, uint32_t bar_cluster_addr, void* dst, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
               ::"r"(smem_u32(dst)), "l"((uint64_t)map), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tc_commit_2sm(uint64_t* bar) {       // arrives on the barrier at this offset in BOTH CTAs
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
               ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
               ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}

// a.tiles_m counts 512-row pair tiles, a.tiles_n 256-column tiles, a.nkb k-blocks of 16; grid = 2 x pairs (even)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(P2_THREADS, 1)
gemm_tc2sm_kernel(const __grid_constant__ CUtensorMap map_ah, const __grid_constant__ CUtensorMap map_al,
                  const __grid_constant__ CUtensorMap map_bh, const __grid_constant__ CUtensorMap map_bl,
                  const __grid_constant__ CUtensorMap map_c, const __grid_constant__ TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // 1024-byte aligned, and still known to be SHARED (LDS / STS, not generic LD / ST)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + P2_STAGES * P2_STAGE + EPI_BYTES);
  uint64_t* full = bars;                       // [P2_STAGES]  the leader's are used: its arrival + both CTAs' bytes
  uint64_t* empty = bars + P2_STAGES;          // [P2_STAGES]  per CTA, multicast commit
  uint64_t* acc_full = bars + 2 * P2_STAGES;   // [1]          per CTA, multicast commit
  uint64_t* acc_empty = acc_full + 1;          // [1]          the leader's is used: 2 x P2_EPI_WARPS arrivals
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int total_tiles = a.tiles_m * a.tiles_n * a.batch;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P2_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, 2 * P2_EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                          // the peer's barriers exist before anybody signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      const uint32_t tx_bytes = a.passes != 1 ? P2_STAGE : P2_APLANE + P2_BHALF;
      const uint32_t full0 = mapa_u32(smem_u32(&full[0]), 0);      // the leader's full[0] in the cluster window
      int s = 0; uint32_t ph = 0;
      for (int t = pair; t < total_tiles; t += npairs) {
        int b, mt, nt;
        tile_coords(t, a, b, mt, nt);
        const int ba = a.a_bcast ? 0 : b, bb = a.b_bcast ? 0 : b;
        const int row_a = mt * 512 + (int)rank * 256, row_b = nt * 256 + (int)rank * 128;
        for (int kb = 0; kb < a.nkb; ++kb) {
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* st = smem + s * P2_STAGE;
          const uint32_t fb = full0 + (uint32_t)s * 8u;
          // ONE arrival per phase, by the leader, for the bytes of both CTAs; the peer only sends bytes.  (Its bytes may land
          // before the leader's expect_tx: the count goes negative inside the same phase, which the pending arrival keeps open.
          // A remote arrive.expect_tx per stage from the peer cost the pipeline half its rate: 674 vs 484 us at 4096^3.)
          if (rank == 0) mbar_expect_tx(&full[s], 2 * tx_bytes);
          tma_load_3d_2sm(&map_ah, fb, st, kb * T2K, row_a, ba);
          tma_load_3d_2sm(&map_bh, fb, st + 2 * P2_APLANE, kb * T2K, row_b, bb);
          if (a.passes != 1) {
            tma_load_3d_2sm(&map_al, fb, st + P2_APLANE, kb * T2K, row_a, ba);
            tma_load_3d_2sm(&map_bl, fb, st + 2 * P2_APLANE + P2_BHALF, kb * T2K, row_b, bb);
          }
          if (++s == P2_STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && elect_one()) {
      int s = 0; uint32_t ph = 0;
      uint32_t aph = 0;
      for (int t = pair; t < total_tiles; t += npairs) {
        mbar_wait(acc_empty, aph ^ 1);          // both CTAs' epilogues have drained both accumulators
        tc_fence_after();
        for (int kb = 0; kb < a.nkb; ++kb) {
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t st = smem_u32(smem + s * P2_STAGE);
          const uint64_t dbh = make_desc_sw64(st + 2 * P2_APLANE), dbl = make_desc_sw64(st + 2 * P2_APLANE + P2_BHALF);
#pragma unroll
          for (int h = 0; h < 2; ++h) {         // rows [0,128) and [128,256) of EACH CTA's A tile
            const uint64_t dah = make_desc_sw64(st + h * (128 * 64)), dal = make_desc_sw64(st + P2_APLANE + h * (128 * 64));
            const uint32_t d = tmem_base + (uint32_t)(h * 256);
            if (a.passes == 2) {
              tc_mma_bf16_2sm(d, dal, dbl + 2, kIdescBf16P2, kb != 0);
              tc_mma_bf16_2sm(d, dal + 2, dbl, kIdescBf16P2, 1);
            }
#pragma unroll
            for (int k = 0; k < T2K / UK; ++k) {
              const uint64_t adv = (uint64_t)((k * UK * 4) >> 4);
              if (a.passes == 2) {
                tc_mma_tf32_2sm(d, dah + adv, dbh + adv, kIdescTf32P2, 1);
              } else if (a.passes == 3) {
                tc_mma_tf32_2sm(d, dal + adv, dbh + adv, kIdescTf32P2, (kb | k) != 0);
                tc_mma_tf32_2sm(d, dah + adv, dbl + adv, kIdescTf32P2, 1);
                tc_mma_tf32_2sm(d, dah + adv, dbh + adv, kIdescTf32P2, 1);
              } else {
                tc_mma_tf32_2sm(d, dah + adv, dbh + adv, kIdescTf32P2, (kb | k) != 0);
              }
            }
          }
          tc_commit_2sm(&empty[s]);
          if (++s == P2_STAGES) { s = 0; ph ^= 1; }
        }
        tc_commit_2sm(acc_full);
        aph ^= 1;
      }
    }
  } else {
    const int q = warp & 3, cgrp = (warp - 2) >> 2;       // TMEM lane quarter (fixed by warp % 4), column half of the tile
    uint8_t* blk = smem + P2_STAGES * P2_STAGE + (warp - 2) * P2_EPI_SLOT;
    const uint32_t acc_empty_leader = mapa_u32(smem_u32(acc_empty), 0);
    uint32_t aph = 0;
    for (int t = pair; t < total_tiles; t += npairs) {
      int b, mt, nt;
      tile_coords(t, a, b, mt, nt);
      mbar_wait(acc_full, aph);
      tc_fence_after();
      float* cbase = a.C + (int64_t)b * a.sc;
      const int row_base = mt * 512 + (int)rank * 256;
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * 256 + cgrp * 128);
        const int row0 = row_base + h * 128 + q * 32;
        uint32_t v0[32], v1[32];
        tc_ld32(tbase, v0);
#pragma unroll 1
        for (int c = 0; c < 128 / 32; c += 2) {
          tc_ld_wait();
          tc_ld32(tbase + (c + 1) * 32, v1);
          int col0 = nt * 256 + cgrp * 128 + c * 32;
          if (col0 < a.n) tc_epilogue32(a, v0, row0, col0, lane);
          if (a.c_tma) { tc_store32_tma1(a, &map_c, blk, v0, row0, col0, b, lane); tc_plane32(a, blk, v0, row0, col0, b, lane); }
          else tc_store32_rows(a, cbase, v0, row0, col0, lane);
          tc_ld_wait();
          if (c + 2 < 128 / 32) tc_ld32(tbase + (c + 2) * 32, v0);
          col0 += 32;
          if (col0 < a.n) tc_epilogue32(a, v1, row0, col0, lane);
          if (a.c_tma) { tc_store32_tma1(a, &map_c, blk, v1, row0, col0, b, lane); tc_plane32(a, blk, v1, row0, col0, b, lane); }
          else tc_store32_rows(a, cbase, v1, row0, col0, lane);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(acc_empty_leader);
      aph ^= 1;
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                          // nobody leaves while the peer may still signal it or read its shared memory
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

}  // namespace wf2
