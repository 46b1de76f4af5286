// ts_mma_test.cu — micro test: tcgen05.mma kind::tf32 with the A operand in TENSOR MEMORY (TS mode).
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -I cl-waffe2_b200/csrc -o /tmp/ts_mma scripts/micro/ts_mma_test.cu
// Checks the layout assumption used by csrc/attention.cu: row r of a 128-row A tile lives in TMEM lane r, element k of the
// row in column (base + k), written by tcgen05.st.32x32b; one MMA consumes 8 columns (K = 8 for tf32).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "tc.cuh"
using namespace wf2;
namespace wf2 { int set_error(int, const char*, ...) { return 1; } EncodeTiledFn tensor_map_encoder() { return nullptr; } float* tc_workspace(size_t) { return nullptr; } }

constexpr int M = 128, N = 32, K = 32;
__device__ __forceinline__ uint32_t sw128(int r, int k) { return (uint32_t)(r * 128 + ((((k >> 2) ^ (r & 7)) << 4) | ((k & 3) << 2))); }
__device__ __forceinline__ void tc_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
               "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
               ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
                 "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]),
                 "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]),
                 "r"(v[30]), "r"(v[31]) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n"
               ::"r"(d), "r"(a_tmem), "l"(db), "r"(idesc), "r"(acc) : "memory");
}

__global__ void __launch_bounds__(128) ts_kernel(const float* A, const float* B, float* D) {
  __shared__ __align__(1024) uint8_t bt[N * 128];
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "n"(128) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = tslot;
  // B (N x K, K-major) into the SWIZZLE_128B layout
  for (int e = threadIdx.x; e < N * K; e += 128) { const int n = e / K, k = e % K; *reinterpret_cast<float*>(bt + sw128(n, k)) = B[n * K + k]; }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  // A row (threadIdx.x) into TMEM columns 64..95 of lane threadIdx.x
  uint32_t v[32];
  for (int k = 0; k < 32; ++k) v[k] = __float_as_uint(A[threadIdx.x * K + k]);
  tc_st32(tb + ((uint32_t)(warp * 32) << 16) + 64, v);
  tc_st_wait();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  if (threadIdx.x == 0) {
    const uint64_t db = make_desc_sw128(smem_u32(bt));
    for (int k = 0; k < K / 8; ++k) mma_ts(tb, tb + 64 + k * 8, db + (uint64_t)((k * 32) >> 4), idesc_tf32_mn(M, N), k != 0);
    tc_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  uint32_t d[32];
  tc_ld32(tb + ((uint32_t)(warp * 32) << 16), d);
  tc_ld_wait();
  for (int n = 0; n < N; ++n) D[threadIdx.x * N + n] = __uint_as_float(d[n]);
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "n"(128) : "memory"); }
}

int main() {
  std::vector<float> A(M * K), B(N * K), D(M * N);
  uint32_t s = 1; auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (float)((s >> 8) & 0xFFFFFF) / 16777216.0f * 2 - 1; };
  for (auto& x : A) x = rnd(); for (auto& x : B) x = rnd();
  float *dA, *dB, *dD;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, D.size() * 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
  ts_kernel<<<1, 128>>>(dA, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("TS-mode MMA: CUDA error %s\n", cudaGetErrorString(e)); return 1; }
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  double worst = 0, mx = 0;
  for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
    double r = 0; for (int k = 0; k < K; ++k) r += (double)A[m * K + k] * B[n * K + k];
    worst = fmax(worst, fabs(r - D[m * N + n])); mx = fmax(mx, fabs(r));
  }
  printf("TS-mode MMA (A in TMEM, lane = row, column = k): max|err| / max|D| = %.3e (one TF32 pass expects ~1e-3; O(1) = wrong layout)\n", worst / mx);
  return 0;
}
