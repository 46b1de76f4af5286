// mma_rate_test.cu — micro benchmark: cycles per tcgen05.mma kind::tf32 (M = 128, K = 8) as a function of N, of the A operand's
// home (shared memory / tensor memory) and of how many independent accumulators the issue stream rotates through.
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -I cl-waffe2_b200/csrc -o /tmp/mma_rate scripts/micro/mma_rate_test.cu
// One CTA, one issuing thread, operands are whatever is in shared / tensor memory (the values do not matter for timing).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include "tc.cuh"
using namespace wf2;
namespace wf2 { int set_error(int, const char*, ...) { return 1; } EncodeTiledFn tensor_map_encoder() { return nullptr; } float* tc_workspace(size_t) { return nullptr; } }

__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n"
               ::"r"(d), "r"(a_tmem), "l"(db), "r"(idesc), "r"(acc) : "memory");
}

// N: MMA N; NACC: accumulators rotated through (each N columns wide); TS: A from TMEM; iters MMAs in total.
// The issue loop is unrolled 16 x with compile-time operands: the first version of this test computed `i % nacc` per MMA and
// measured its own integer division (188 cycles per iteration whatever N was).
template <int N, int NACC, int TS, int ELECT>
__global__ void __launch_bounds__(128) rate_kernel(int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];      // A tile 128 x 32 (16 KiB) | B tile 256 x 32 (32 KiB)
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 128) reinterpret_cast<float*>(smem)[i] = 1.0f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = tslot;
  // ELECT=1: the issuing branch is taken under elect.sync (the compiler then emits the UTCHMMAs back to back); 0: `threadIdx.x == 0`
  // (every UTCHMMA wrapped in an ELECT / BRA.U.ANY loop over the active lanes: what the first version of this test measured)
  if (ELECT ? (warp == 0 && elect_one()) : (threadIdx.x == 0)) {
    const uint64_t da = make_desc_sw128(smem_u32(smem)), db = make_desc_sw128(smem_u32(smem + 16384));
    constexpr uint32_t idesc = idesc_tf32_mn(128, N);
    constexpr uint32_t acol = 448;                        // A operand columns when TS (accumulators use < 448)
    long long t0 = clock64();
    for (int i = 0; i < iters; i += 16) {
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const uint32_t d = tb + (uint32_t)((u % NACC) * N);
        const uint64_t adv = (uint64_t)(((u & 3) * 32) >> 4);
        if (TS) mma_ts(d, tb + acol + (u & 3) * 8, db + adv, idesc, 1);
        else tc_mma_tf32(d, da + adv, db + adv, idesc, 1);
      }
    }
    tc_commit(&bar);
    mbar_wait(&bar, 0);
    long long t1 = clock64();
    out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "n"(512) : "memory"); }
}

template <int N, int NACC>
static void run(long long* d, int iters) {
  double c[4];
  for (int v = 0; v < 4; ++v) {
    for (int rep = 0; rep < 2; ++rep) {
      if (v == 0) rate_kernel<N, NACC, 0, 0><<<1, 128, 49152 + 1024>>>(iters, d);
      if (v == 1) rate_kernel<N, NACC, 1, 0><<<1, 128, 49152 + 1024>>>(iters, d);
      if (v == 2) rate_kernel<N, NACC, 0, 1><<<1, 128, 49152 + 1024>>>(iters, d);
      if (v == 3) rate_kernel<N, NACC, 1, 1><<<1, 128, 49152 + 1024>>>(iters, d);
    }
    long long h; cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); exit(1); }
    c[v] = (double)h / iters;
  }
  printf("%6d %6d %10.1f %10.1f %12.1f %12.1f   (%d)\n", N, NACC, c[0], c[1], c[2], c[3], 128 * N / 256);
}

int main() {
  long long* d; cudaMalloc(&d, 8);
#define ATTR(N, A) cudaFuncSetAttribute(rate_kernel<N, A, 0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152 + 1024); \
                   cudaFuncSetAttribute(rate_kernel<N, A, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152 + 1024); \
                   cudaFuncSetAttribute(rate_kernel<N, A, 0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152 + 1024); \
                   cudaFuncSetAttribute(rate_kernel<N, A, 1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152 + 1024);
  ATTR(16, 1) ATTR(16, 4) ATTR(32, 1) ATTR(32, 4) ATTR(64, 1) ATTR(64, 4) ATTR(128, 1) ATTR(128, 2) ATTR(256, 1)
  const int iters = 4096;
  printf("cycles per tcgen05.mma.kind::tf32 128 x N x 8 (one issuing thread, %d MMAs back to back, unrolled issue)\n", iters);
  printf("%6s %6s %10s %10s %12s %12s   (128 N / 256)\n", "N", "nacc", "A in smem", "A in tmem", "smem, elect", "tmem, elect");
  run<16, 1>(d, iters); run<16, 4>(d, iters); run<32, 1>(d, iters); run<32, 4>(d, iters); run<64, 1>(d, iters); run<64, 4>(d, iters);
  run<128, 1>(d, iters); run<128, 2>(d, iters); run<256, 1>(d, iters);
  return 0;
}
