// mma_rate_2sm_test.cu — cycles per tcgen05.mma.cta_group::2 (M = 256 over a CTA pair, N = 256) for kind::tf32 (K = 8) and
// kind::f16 / bf16 (K = 16), against the single-CTA instruction (M = 128, N = 256) on the same data.
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -I cl-waffe2_b200/csrc -o /tmp/mma_rate_2sm scripts/micro/mma_rate_2sm_test.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include "tc.cuh"
using namespace wf2;
namespace wf2 { int set_error(int, const char*, ...) { return 1; } EncodeTiledFn tensor_map_encoder() { return nullptr; } float* tc_workspace(size_t) { return nullptr; } }

__device__ __forceinline__ uint32_t ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void csync() { asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }
template <int KIND>   // 0 tf32, 1 bf16
__device__ __forceinline__ void mma2(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc) {
  if (KIND == 0) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d), "l"(da), "l"(db), "r"(idesc) : "memory");
  else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d), "l"(da), "l"(db), "r"(idesc) : "memory");
}
template <int KIND>
__device__ __forceinline__ void mma1(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc) {
  if (KIND == 0) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d), "l"(da), "l"(db), "r"(idesc) : "memory");
  else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d), "l"(da), "l"(db), "r"(idesc) : "memory");
}
__host__ __device__ constexpr uint32_t idesc_bf16(int m, int n) { return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24); }

// PAIR = 1: cluster of 2, leader issues cta_group::2 MMAs (B: 128 rows per CTA); PAIR = 0: every CTA issues its own cta_group::1 MMAs
template <int PAIR, int KIND>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128) rate2_kernel(int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];      // A 128 x 128 B (16 KiB) | B 256 x 128 B (32 KiB)
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = ctarank();
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 128) reinterpret_cast<float*>(smem)[i] = 1.0f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    if (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "n"(512) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "n"(512) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before(); __syncthreads(); csync(); tc_fence_after();
  const uint32_t tb = tslot;
  if (threadIdx.x == 0 && (!PAIR || rank == 0)) {
    const uint64_t da = make_desc_sw128(smem_u32(smem)), db = make_desc_sw128(smem_u32(smem + 16384));
    const uint32_t idesc = KIND == 0 ? idesc_tf32_mn(PAIR ? 256 : 128, 256) : idesc_bf16(PAIR ? 256 : 128, 256);
    long long t0 = clock64();
    for (int i = 0; i < iters; i += 16) {
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const uint32_t d = tb + (uint32_t)((u & 1) * 256);
        const uint64_t adv = (uint64_t)(((u & 3) * 32) >> 4);
        if (PAIR) mma2<KIND>(d, da + adv, db + adv, idesc);
        else mma1<KIND>(d, da + adv, db + adv, idesc);
      }
    }
    if (PAIR) asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "h"((uint16_t)3) : "memory");
    else tc_commit(&bar);
    mbar_wait(&bar, 0);
    long long t1 = clock64();
    out[rank] = t1 - t0;
  } else if (PAIR && threadIdx.x == 0) {
    mbar_wait(&bar, 0);
  }
  tc_fence_before(); __syncthreads(); csync();
  if (warp == 0) {
    tc_fence_after();
    if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tb), "n"(512) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "n"(512) : "memory");
  }
}

template <int PAIR, int KIND>
static void run(long long* d, int iters, int grid, const char* what) {
  cudaFuncSetAttribute(rate2_kernel<PAIR, KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152 + 1024);
  for (int rep = 0; rep < 2; ++rep) rate2_kernel<PAIR, KIND><<<grid, 128, 49152 + 1024>>>(iters, d);
  long long h[2]; cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { printf("%s: CUDA error %s\n", what, cudaGetErrorString(e)); exit(1); }
  printf("%-64s %8.1f cycles per MMA (grid %d)\n", what, (double)h[0] / iters, grid);
}

int main() {
  long long* d; cudaMalloc(&d, 16);
  const int iters = 4096;
  for (int grid : {2, 148}) {
    run<0, 0>(d, iters, grid, "cta_group::1 kind::tf32 128 x 256 x 8  (per CTA)");
    run<1, 0>(d, iters, grid, "cta_group::2 kind::tf32 256 x 256 x 8  (per pair)");
    run<0, 1>(d, iters, grid, "cta_group::1 kind::f16 (bf16) 128 x 256 x 16 (per CTA)");
    run<1, 1>(d, iters, grid, "cta_group::2 kind::f16 (bf16) 256 x 256 x 16 (per pair)");
  }
  return 0;
}
