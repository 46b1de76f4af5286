// redbench.cu — micro-benchmark (NOT product code): strategies for a 64 MiB read-only fp32 sum on B200.
// Build: python scripts/mk_redbench_product.py && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/redbench scripts/redbench.cu
// Each variant is timed as 20 launches (rotating over 4 x 64 MiB buffers, > L2) inside one CUDA graph.
#include <cuda_runtime.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cstring>
#include <cmath>
static const char* g_filter = nullptr;
constexpr int kMaxRank = 8;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ float warp_sum(float v) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <int T>
__device__ __forceinline__ float block_sum(float v, float* sm) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < T / 32 ? sm[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) sm[0] = t;
  }
  __syncthreads();
  return sm[0];
}
__device__ __forceinline__ float sum4(float4 a) { return (a.x + a.y) + (a.z + a.w); }

// ---- V0: empty kernel (launch floor)
__global__ void k_fill(float* p, size_t n, uint32_t seed) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    uint32_t h = (uint32_t)i * 2654435761u + seed; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    p[i] = (float)(h & 0xFFFFFF) * (2.0f / 16777216.0f) - 1.0f;
  }
}
__global__ void k_empty(const float4* p, float* out, size_t n4) {}

// ---- V1: grid-stride, U float4 loads in flight per thread, partials per block
template <int T, int U>
__global__ void __launch_bounds__(T) k_stride(const float4* __restrict__ p, float* __restrict__ out, size_t n4) {
  __shared__ float sm[32];
  float acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = 0.f;
  size_t i = (size_t)blockIdx.x * T * U + threadIdx.x;
  const size_t step = (size_t)gridDim.x * T * U;
  for (; i + (size_t)(U - 1) * T < n4; i += step) {
    float4 x[U];
#pragma unroll
    for (int k = 0; k < U; ++k) x[k] = p[i + (size_t)k * T];
#pragma unroll
    for (int k = 0; k < U; ++k) acc[k] += sum4(x[k]);
  }
  float v = 0.f;
#pragma unroll
  for (int k = 0; k < U; ++k) v += acc[k];
  v = block_sum<T>(v, sm);
  if (threadIdx.x == 0) out[blockIdx.x] = v;
}

// ---- V2: same with ld.global.nc.L1::no_allocate.L2::256B hint
__device__ __forceinline__ float4 ldnc(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::256B.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
template <int T, int U>
__global__ void __launch_bounds__(T) k_stride_nc(const float4* __restrict__ p, float* __restrict__ out, size_t n4) {
  __shared__ float sm[32];
  float acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = 0.f;
  size_t i = (size_t)blockIdx.x * T * U + threadIdx.x;
  const size_t step = (size_t)gridDim.x * T * U;
  for (; i + (size_t)(U - 1) * T < n4; i += step) {
    float4 x[U];
#pragma unroll
    for (int k = 0; k < U; ++k) x[k] = ldnc(p + i + (size_t)k * T);
#pragma unroll
    for (int k = 0; k < U; ++k) acc[k] += sum4(x[k]);
  }
  float v = 0.f;
#pragma unroll
  for (int k = 0; k < U; ++k) v += acc[k];
  v = block_sum<T>(v, sm);
  if (threadIdx.x == 0) out[blockIdx.x] = v;
}

// ---- V3: bulk async copy (TMA 1-D) into a shared-memory ring, persistent CTAs
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int T, int STAGES, int CHUNK>   // CHUNK bytes per stage
__global__ void __launch_bounds__(T) k_bulk(const float4* __restrict__ p, float* __restrict__ out, size_t n4) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar[STAGES];
  __shared__ float sm[32];
  const size_t total_bytes = n4 * 16;
  const size_t nchunks = total_bytes / CHUNK;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const char* base = reinterpret_cast<const char*>(p);
  // prologue
  size_t next = blockIdx.x;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES && next < nchunks; ++s, next += gridDim.x) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar[s])), "r"(CHUNK) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(s32(smem + s * CHUNK)), "l"(base + next * CHUNK), "r"(CHUNK), "r"(s32(&bar[s])) : "memory");
    }
  }
  float acc0 = 0.f, acc1 = 0.f;
  int s = 0; uint32_t ph = 0;
  for (size_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
    uint32_t ok = 0;
    while (!ok) {
      asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\tselp.u32 %0,1,0,q;\n\t}" : "=r"(ok) : "r"(s32(&bar[s])), "r"(ph) : "memory");
    }
    const float4* t = reinterpret_cast<const float4*>(smem + s * CHUNK);
#pragma unroll 4
    for (int i = threadIdx.x; i < CHUNK / 16; i += 2 * T) {
      acc0 += sum4(t[i]);
      if (i + T < CHUNK / 16) acc1 += sum4(t[i + T]);
    }
    __syncthreads();     // everyone done with stage s
    if (threadIdx.x == 0) {
      const size_t nc = c + (size_t)STAGES * gridDim.x;
      if (nc < nchunks) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar[s])), "r"(CHUNK) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(s32(smem + s * CHUNK)), "l"(base + nc * CHUNK), "r"(CHUNK), "r"(s32(&bar[s])) : "memory");
      }
    }
    if (++s == STAGES) { s = 0; ph ^= 1; }
  }
  float v = block_sum<T>(acc0 + acc1, sm);
  if (threadIdx.x == 0) out[blockIdx.x] = v;
}


// ---- V4: single launch, last-block ticket.  FENCE=1: __threadfence + atomicAdd (the product's scheme);
//          FENCE=0: one atom.acq_rel.gpu (release of the partial store, acquire for the last block)
template <int T, int U, int FENCE>
__global__ void __launch_bounds__(T) k_ticket(const float4* __restrict__ p, float* __restrict__ part, float* __restrict__ out,
                                              unsigned int* counter, size_t n4) {
  __shared__ float sm[32];
  __shared__ bool last;
  float acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = 0.f;
  size_t i = (size_t)blockIdx.x * T * U + threadIdx.x;
  const size_t step = (size_t)gridDim.x * T * U;
  for (; i + (size_t)(U - 1) * T < n4; i += step) {
    float4 x[U];
#pragma unroll
    for (int k = 0; k < U; ++k) x[k] = p[i + (size_t)k * T];
#pragma unroll
    for (int k = 0; k < U; ++k) acc[k] += sum4(x[k]);
  }
  float v = 0.f;
#pragma unroll
  for (int k = 0; k < U; ++k) v += acc[k];
  v = block_sum<T>(v, sm);
  if (threadIdx.x == 0) {
    unsigned int t;
    if (FENCE) {
      part[blockIdx.x] = v;
      __threadfence();
      t = atomicAdd(counter, 1u);
    } else {
      asm volatile("st.relaxed.gpu.global.f32 [%0], %1;" ::"l"(part + blockIdx.x), "f"(v) : "memory");
      asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(counter) : "memory");
    }
    last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!last) return;
  if (FENCE) __threadfence();
  float w = 0.f;
  for (int c = threadIdx.x; c < gridDim.x; c += T) w += __ldcg(part + c);
  w = block_sum<T>(w, sm);
  if (threadIdx.x == 0) { out[0] = w; *counter = 0; }
}

// ---- rows (sum :axis 1 of a 4096 x 4096 matrix): one-shot tile kernel as in the product (RPC rows per CTA)
template <int RPC>
__global__ void __launch_bounds__(256) k_rows_oneshot(const float4* __restrict__ p, float* __restrict__ out, int rows, int v1) {
  constexpr int TPR = 256 / RPC, WPR = TPR / 32;
  __shared__ float sm[8];
  const int sub = threadIdx.x / TPR, t = threadIdx.x % TPR;
  const int row = blockIdx.x * RPC + sub;
  const float4* q = p + (size_t)row * v1;
  float4 x[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) x[k] = q[t + k * TPR];
  float v = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) v += sum4(x[k]);
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  if (t == 0) { float r = 0.f; for (int k = 0; k < WPR; ++k) r += sm[sub * WPR + k]; out[row] = r; }
}
// ---- rows: a CTA per row, U4 per thread (row = 1024 float4), looping over G rows with register prefetch
__global__ void __launch_bounds__(256) k_rows_loop(const float4* __restrict__ p, float* __restrict__ out, int rows, int v1, int G) {
  __shared__ float sm[2][8];
  const int r0 = blockIdx.x * G;
  float4 x[4], y[4];
  const float4* q = p + (size_t)r0 * v1;
#pragma unroll
  for (int k = 0; k < 4; ++k) x[k] = q[threadIdx.x + k * 256];
  for (int g = 0; g < G; ++g) {
    if (g + 1 < G) {
      const float4* qn = p + (size_t)(r0 + g + 1) * v1;
#pragma unroll
      for (int k = 0; k < 4; ++k) y[k] = qn[threadIdx.x + k * 256];
    }
    float v = (sum4(x[0]) + sum4(x[1])) + (sum4(x[2]) + sum4(x[3]));
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sm[g & 1][threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) { float r = 0.f; for (int k = 0; k < 8; ++k) r += sm[g & 1][k]; out[r0 + g] = r; }
#pragma unroll
    for (int k = 0; k < 4; ++k) x[k] = y[k];
  }
}
// ---- rows: a WARP per row group, no block sync; each lane 32 float4 per row in 4 batches of 8
__global__ void __launch_bounds__(256) k_rows_warp(const float4* __restrict__ p, float* __restrict__ out, int rows, int v1, int G) {
  const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  for (int g = 0; g < G; ++g) {
    const int row = warp * G + g;
    if (row >= rows) return;
    const float4* q = p + (size_t)row * v1;
    float a0 = 0.f, a1 = 0.f;
    for (int i = lane; i < v1; i += 256) {
      float4 x[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) x[k] = q[i + k * 32];
#pragma unroll
      for (int k = 0; k < 8; k += 2) { a0 += sum4(x[k]); a1 += sum4(x[k + 1]); }
    }
    const float v = warp_sum(a0 + a1);
    if (lane == 0) out[row] = v;
  }
}
// ---- cols (sum :axis 0): thread owns 4 columns (float4), CTA = 32 x 8, grid (cols/128, splits); partial rows out
template <int U>
__global__ void __launch_bounds__(256) k_cols(const float4* __restrict__ p, float4* __restrict__ part, int rows, int v1, int rows_per_split) {
  __shared__ float4 sm[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  const int r0 = blockIdx.y * rows_per_split, r1 = min(rows, r0 + rows_per_split);
  float4 acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = make_float4(0, 0, 0, 0);
  int r = r0 + ty;
  for (; r + 8 * (U - 1) < r1; r += 8 * U) {
    float4 x[U];
#pragma unroll
    for (int k = 0; k < U; ++k) x[k] = p[(size_t)(r + 8 * k) * v1 + c];
#pragma unroll
    for (int k = 0; k < U; ++k) { acc[k].x += x[k].x; acc[k].y += x[k].y; acc[k].z += x[k].z; acc[k].w += x[k].w; }
  }
  for (; r < r1; r += 8) { const float4 x = p[(size_t)r * v1 + c]; acc[0].x += x.x; acc[0].y += x.y; acc[0].z += x.z; acc[0].w += x.w; }
  float4 a = acc[0];
#pragma unroll
  for (int k = 1; k < U; ++k) { a.x += acc[k].x; a.y += acc[k].y; a.z += acc[k].z; a.w += acc[k].w; }
  sm[ty][tx] = a;
  __syncthreads();
  if (ty == 0) {
    for (int k = 1; k < 8; ++k) { const float4 b = sm[k][tx]; a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
    part[(size_t)blockIdx.y * v1 + c] = a;
  }
}
// cols with a per-column-block ticket: the last split folds the partial rows (single launch)
template <int U>
__global__ void __launch_bounds__(256) k_cols_ticket(const float4* __restrict__ p, float4* __restrict__ part, float4* __restrict__ out,
                                                      unsigned int* counters, int rows, int v1, int rows_per_split) {
  __shared__ float4 sm[8][32];
  __shared__ bool last;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  const int r0 = blockIdx.y * rows_per_split, r1 = min(rows, r0 + rows_per_split);
  float4 acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = make_float4(0, 0, 0, 0);
  int r = r0 + ty;
  for (; r + 8 * (U - 1) < r1; r += 8 * U) {
    float4 x[U];
#pragma unroll
    for (int k = 0; k < U; ++k) x[k] = p[(size_t)(r + 8 * k) * v1 + c];
#pragma unroll
    for (int k = 0; k < U; ++k) { acc[k].x += x[k].x; acc[k].y += x[k].y; acc[k].z += x[k].z; acc[k].w += x[k].w; }
  }
  for (; r < r1; r += 8) { const float4 x = p[(size_t)r * v1 + c]; acc[0].x += x.x; acc[0].y += x.y; acc[0].z += x.z; acc[0].w += x.w; }
  float4 a = acc[0];
#pragma unroll
  for (int k = 1; k < U; ++k) { a.x += acc[k].x; a.y += acc[k].y; a.z += acc[k].z; a.w += acc[k].w; }
  sm[ty][tx] = a;
  __syncthreads();
  if (ty == 0) {
    for (int k = 1; k < 8; ++k) { const float4 b = sm[k][tx]; a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
    part[(size_t)blockIdx.y * v1 + c] = a;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) { const unsigned int t = atomicAdd(&counters[blockIdx.x], 1u); last = (t == gridDim.y - 1); }
  __syncthreads();
  if (!last) return;
  __threadfence();
  float4 t4 = make_float4(0, 0, 0, 0);
  for (int s = ty; s < (int)gridDim.y; s += 8) { const float4 b = __ldcg(part + (size_t)s * v1 + c); t4.x += b.x; t4.y += b.y; t4.z += b.z; t4.w += b.w; }
  __syncthreads();
  sm[ty][tx] = t4;
  __syncthreads();
  if (ty == 0) {
    for (int k = 1; k < 8; ++k) { const float4 b = sm[k][tx]; t4.x += b.x; t4.y += b.y; t4.z += b.z; t4.w += b.w; }
    out[c] = t4;
    if (tx == 0) counters[blockIdx.x] = 0;
  }
}
// cols with an 8-CTA cluster along the reduced axis: the 8 partial column strips meet in distributed shared
// memory, rank 0 folds them and writes the result -- no partials in global memory, no ticket, one launch.
template <int U, int LX>
__global__ void __cluster_dims__(1, 8, 1) __launch_bounds__(256) k_cols_cluster(const float4* __restrict__ p, float4* __restrict__ out, int rows, int v1, int rows_per_split) {
  constexpr int LY = 256 / LX;
  __shared__ float4 sm[LY][LX];
  __shared__ float4 res[LX];
  cg::cluster_group cluster = cg::this_cluster();
  const int tx = threadIdx.x % LX, ty = threadIdx.x / LX;
  const int c = blockIdx.x * LX + tx;
  const int r0 = blockIdx.y * rows_per_split, r1 = min(rows, r0 + rows_per_split);
  float4 acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = make_float4(0, 0, 0, 0);
  int r = r0 + ty;
  for (; r + LY * (U - 1) < r1; r += LY * U) {
    float4 x[U];
#pragma unroll
    for (int k = 0; k < U; ++k) x[k] = p[(size_t)(r + LY * k) * v1 + c];
#pragma unroll
    for (int k = 0; k < U; ++k) { acc[k].x += x[k].x; acc[k].y += x[k].y; acc[k].z += x[k].z; acc[k].w += x[k].w; }
  }
  for (; r < r1; r += LY) { const float4 x = p[(size_t)r * v1 + c]; acc[0].x += x.x; acc[0].y += x.y; acc[0].z += x.z; acc[0].w += x.w; }
  float4 a = acc[0];
#pragma unroll
  for (int k = 1; k < U; ++k) { a.x += acc[k].x; a.y += acc[k].y; a.z += acc[k].z; a.w += acc[k].w; }
  sm[ty][tx] = a;
  __syncthreads();
  if (ty == 0) {
#pragma unroll
    for (int k = 1; k < LY; ++k) { const float4 b = sm[k][tx]; a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
    res[tx] = a;
  }
  cluster.sync();
  if (cluster.block_rank() == 0 && ty == 0) {
    float4 t[8];
#pragma unroll
    for (int k = 1; k < 8; ++k) t[k] = *cluster.map_shared_rank(&res[tx], k);
#pragma unroll
    for (int k = 1; k < 8; ++k) { a.x += t[k].x; a.y += t[k].y; a.z += t[k].z; a.w += t[k].w; }
    out[c] = a;
  }
  cluster.sync();
}
__global__ void k_cols_final(const float4* part, float4* out, int splits, int v1) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= v1) return;
  float4 a = make_float4(0, 0, 0, 0);
  for (int s = 0; s < splits; ++s) { const float4 b = part[(size_t)s * v1 + c]; a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
  out[c] = a;
}

// ---- final: fold partials (second launch) — measures the cost of a 2-launch scheme
__global__ void k_final(const float* part, int n, float* out) {
  __shared__ float sm[32];
  float v = 0.f;
  for (int i = threadIdx.x; i < n; i += 256) v += part[i];
  v = block_sum<256>(v, sm);
  if (threadIdx.x == 0) out[0] = v;
}

#include "redbench_product.cuh"
template <class F>
static int time_variant(const char* name, cudaStream_t st, F launch, size_t bytes) {
  if (g_filter && !strstr(name, g_filter)) return 0;
  for (int i = 0; i < 3; ++i) launch(i & 3);
  CK(cudaStreamSynchronize(st));
  CK(cudaGetLastError());
  cudaGraph_t g; cudaGraphExec_t ex;
  CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
  for (int i = 0; i < 20; ++i) launch(i & 3);
  CK(cudaStreamEndCapture(st, &g));
  CK(cudaGraphInstantiate(&ex, g, 0));
  CK(cudaGraphLaunch(ex, st));
  CK(cudaStreamSynchronize(st));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0, st); cudaGraphLaunch(ex, st); cudaEventRecord(e1, st); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  const double us = best * 1e3 / 20;
  printf("%-34s %8.2f us  %8.1f GB/s\n", name, us, bytes / us * 1e-3);
  fflush(stdout);
  cudaGraphExecDestroy(ex); cudaGraphDestroy(g);

  return 0;
}

int main(int argc, char** argv) {
  if (argc > 1) g_filter = argv[1];
  const size_t n = (size_t)4096 * 4096, n4 = n / 4, bytes = n * 4;
  float* buf[4]; float* part; float* out;
  for (int i = 0; i < 4; ++i) { CK(cudaMalloc(&buf[i], bytes)); k_fill<<<1184, 256>>>(buf[i], n, 17u * i + 1u); }
  CK(cudaDeviceSynchronize());
  CK(cudaMalloc(&part, 4 << 20)); CK(cudaMalloc(&out, 64));
  float* out2; CK(cudaMalloc(&out2, 1 << 16));
  cudaStream_t st; CK(cudaStreamCreate(&st));
#define P4(i) reinterpret_cast<const float4*>(buf[i])
  time_variant("empty kernel", st, [&](int i) { k_empty<<<148, 256, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T256 U8 grid1024", st, [&](int i) { k_stride<256, 8><<<1024, 256, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T256 U8 grid2048", st, [&](int i) { k_stride<256, 8><<<2048, 256, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T256 U8 grid592", st, [&](int i) { k_stride<256, 8><<<592, 256, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T256 U8 grid1184", st, [&](int i) { k_stride<256, 8><<<1184, 256, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T256 U4 grid1184", st, [&](int i) { k_stride<256, 4><<<1184, 256, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T256 U4 grid2368", st, [&](int i) { k_stride<256, 4><<<2368, 256, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T512 U8 grid592", st, [&](int i) { k_stride<512, 8><<<592, 512, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T512 U4 grid592", st, [&](int i) { k_stride<512, 4><<<592, 512, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T1024 U4 grid296", st, [&](int i) { k_stride<1024, 4><<<296, 1024, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T256 U16 grid592", st, [&](int i) { k_stride<256, 16><<<592, 256, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T128 U16 grid1184", st, [&](int i) { k_stride<128, 16><<<1184, 128, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride.nc T256 U8 grid1184", st, [&](int i) { k_stride_nc<256, 8><<<1184, 256, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride.nc T512 U8 grid592", st, [&](int i) { k_stride_nc<512, 8><<<592, 512, 0, st>>>(P4(i), part, n4); }, bytes);
  time_variant("stride T256 U8 grid1184 + final", st, [&](int i) { k_stride<256, 8><<<1184, 256, 0, st>>>(P4(i), part, n4); k_final<<<1, 256, 0, st>>>(part, 1184, out); }, bytes);
  {
    auto kb = k_bulk<256, 4, 32768>;
    CK(cudaFuncSetAttribute(kb, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 32768));
    time_variant("bulk T256 4x32K grid148", st, [&](int i) { kb<<<148, 256, 4 * 32768, st>>>(P4(i), part, n4); }, bytes);
    auto kb2 = k_bulk<256, 4, 16384>;
    CK(cudaFuncSetAttribute(kb2, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 16384));
    time_variant("bulk T256 4x16K grid296", st, [&](int i) { kb2<<<296, 256, 4 * 16384, st>>>(P4(i), part, n4); }, bytes);
    time_variant("bulk T256 4x16K grid444", st, [&](int i) { kb2<<<444, 256, 4 * 16384, st>>>(P4(i), part, n4); }, bytes);
    auto kb3 = k_bulk<512, 6, 32768>;
    CK(cudaFuncSetAttribute(kb3, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 32768));
    time_variant("bulk T512 6x32K grid148", st, [&](int i) { kb3<<<148, 512, 6 * 32768, st>>>(P4(i), part, n4); }, bytes);
    auto kb4 = k_bulk<256, 8, 8192>;
    CK(cudaFuncSetAttribute(kb4, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 8192));
    time_variant("bulk T256 8x8K grid444", st, [&](int i) { kb4<<<444, 256, 8 * 8192, st>>>(P4(i), part, n4); }, bytes);
  }

  unsigned int* counter; CK(cudaMalloc(&counter, 64)); CK(cudaMemset(counter, 0, 64));
  time_variant("ticket fence  T256 U8 grid1024", st, [&](int i) { k_ticket<256, 8, 1><<<1024, 256, 0, st>>>(P4(i), part, out, counter, n4); }, bytes);
  time_variant("ticket acqrel T256 U8 grid1024", st, [&](int i) { k_ticket<256, 8, 0><<<1024, 256, 0, st>>>(P4(i), part, out, counter, n4); }, bytes);
  time_variant("ticket acqrel T256 U4 grid2368", st, [&](int i) { k_ticket<256, 4, 0><<<2368, 256, 0, st>>>(P4(i), part, out, counter, n4); }, bytes);
  time_variant("ticket acqrel T256 U8 grid1184", st, [&](int i) { k_ticket<256, 8, 0><<<1184, 256, 0, st>>>(P4(i), part, out, counter, n4); }, bytes);
  time_variant("rows oneshot RPC2 grid2048", st, [&](int i) { k_rows_oneshot<2><<<2048, 256, 0, st>>>(P4(i), part, 4096, 1024); }, bytes);
  time_variant("rows loop G4 grid1024", st, [&](int i) { k_rows_loop<<<1024, 256, 0, st>>>(P4(i), part, 4096, 1024, 4); }, bytes);
  time_variant("rows loop G2 grid2048", st, [&](int i) { k_rows_loop<<<2048, 256, 0, st>>>(P4(i), part, 4096, 1024, 2); }, bytes);
  time_variant("rows loop G1 grid4096", st, [&](int i) { k_rows_loop<<<4096, 256, 0, st>>>(P4(i), part, 4096, 1024, 1); }, bytes);
  time_variant("rows warp G1 grid512", st, [&](int i) { k_rows_warp<<<512, 256, 0, st>>>(P4(i), part, 4096, 1024, 1); }, bytes);
  time_variant("rows oneshot RPC2 grid2048 (rnd)", st, [&](int i) { k_rows_oneshot<2><<<2048, 256, 0, st>>>(P4(i), part, 4096, 1024); }, bytes);
  time_variant("cols U4 grid(32,37)", st, [&](int i) { k_cols<4><<<dim3(32, 37), 256, 0, st>>>(P4(i), (float4*)part, 4096, 1024, 111); }, bytes);
  time_variant("cols U8 grid(32,37)", st, [&](int i) { k_cols<8><<<dim3(32, 37), 256, 0, st>>>(P4(i), (float4*)part, 4096, 1024, 111); }, bytes);
  time_variant("cols U4 grid(32,64)", st, [&](int i) { k_cols<4><<<dim3(32, 64), 256, 0, st>>>(P4(i), (float4*)part, 4096, 1024, 64); }, bytes);
  time_variant("cols U8 grid(32,64)", st, [&](int i) { k_cols<8><<<dim3(32, 64), 256, 0, st>>>(P4(i), (float4*)part, 4096, 1024, 64); }, bytes);
  time_variant("cols U4 grid(32,128)", st, [&](int i) { k_cols<4><<<dim3(32, 128), 256, 0, st>>>(P4(i), (float4*)part, 4096, 1024, 32); }, bytes);
  time_variant("cols U8 grid(32,32)", st, [&](int i) { k_cols<8><<<dim3(32, 32), 256, 0, st>>>(P4(i), (float4*)part, 4096, 1024, 128); }, bytes);
  time_variant("cols U8 grid(32,64) + final", st, [&](int i) { k_cols<8><<<dim3(32, 64), 256, 0, st>>>(P4(i), (float4*)part, 4096, 1024, 64); k_cols_final<<<4, 256, 0, st>>>((const float4*)part, (float4*)out2, 64, 1024); }, bytes);
  time_variant("cols ticket U8 grid(32,64)", st, [&](int i) { k_cols_ticket<8><<<dim3(32, 64), 256, 0, st>>>(P4(i), (float4*)part, (float4*)out2, counter, 4096, 1024, 64); }, bytes);
  time_variant("cols ticket U4 grid(32,64)", st, [&](int i) { k_cols_ticket<4><<<dim3(32, 64), 256, 0, st>>>(P4(i), (float4*)part, (float4*)out2, counter, 4096, 1024, 64); }, bytes);
  time_variant("cols ticket U8 grid(32,37)", st, [&](int i) { k_cols_ticket<8><<<dim3(32, 37), 256, 0, st>>>(P4(i), (float4*)part, (float4*)out2, counter, 4096, 1024, 111); }, bytes);
  time_variant("cols ticket U4 grid(32,128)", st, [&](int i) { k_cols_ticket<4><<<dim3(32, 128), 256, 0, st>>>(P4(i), (float4*)part, (float4*)out2, counter, 4096, 1024, 32); }, bytes);
  time_variant("cols cluster8 U8 LX32 grid(32,8)", st, [&](int i) { k_cols_cluster<8, 32><<<dim3(32, 8), 256, 0, st>>>(P4(i), (float4*)out2, 4096, 1024, 512); }, bytes);
  time_variant("cols cluster8 U8 LX16 grid(64,8)", st, [&](int i) { k_cols_cluster<8, 16><<<dim3(64, 8), 256, 0, st>>>(P4(i), (float4*)out2, 4096, 1024, 512); }, bytes);
  time_variant("cols cluster8 U4 LX16 grid(64,8)", st, [&](int i) { k_cols_cluster<4, 16><<<dim3(64, 8), 256, 0, st>>>(P4(i), (float4*)out2, 4096, 1024, 512); }, bytes);
  time_variant("cols cluster8 U8 LX8 grid(128,8)", st, [&](int i) { k_cols_cluster<8, 8><<<dim3(128, 8), 256, 0, st>>>(P4(i), (float4*)out2, 4096, 1024, 512); }, bytes);
  time_variant("cols cluster8 U4 LX8 grid(128,8)", st, [&](int i) { k_cols_cluster<4, 8><<<dim3(128, 8), 256, 0, st>>>(P4(i), (float4*)out2, 4096, 1024, 512); }, bytes);
  time_variant("cols cluster8 U4 LX4 grid(256,8)", st, [&](int i) { k_cols_cluster<4, 4><<<dim3(256, 8), 256, 0, st>>>(P4(i), (float4*)out2, 4096, 1024, 512); }, bytes);
  time_variant("cols ticket U8 grid(32,18)", st, [&](int i) { k_cols_ticket<8><<<dim3(32, 18), 256, 0, st>>>(P4(i), (float4*)part, (float4*)out2, counter, 4096, 1024, 228); }, bytes);
  time_variant("cols ticket U8 grid(32,9)", st, [&](int i) { k_cols_ticket<8><<<dim3(32, 9), 256, 0, st>>>(P4(i), (float4*)part, (float4*)out2, counter, 4096, 1024, 456); }, bytes);
  time_variant("ticket acqrel T256 U8 grid592", st, [&](int i) { k_ticket<256, 8, 0><<<592, 256, 0, st>>>(P4(i), part, out, counter, n4); }, bytes);
  time_variant("ticket acqrel T512 U4 grid592", st, [&](int i) { k_ticket<512, 4, 0><<<592, 512, 0, st>>>(P4(i), part, out, counter, n4); }, bytes);
  time_variant("ticket acqrel T512 U8 grid296", st, [&](int i) { k_ticket<512, 8, 0><<<296, 512, 0, st>>>(P4(i), part, out, counter, n4); }, bytes);
  time_variant("ticket acqrel T256 U8 grid2048", st, [&](int i) { k_ticket<256, 8, 0><<<2048, 256, 0, st>>>(P4(i), part, out, counter, n4); }, bytes);
  {
    using namespace prod;
    RowArgs ra; memset(&ra, 0, sizeof(ra));
    ra.nkept = 1; ra.kshape[0] = 4096; ra.kstride_src[0] = 4096; ra.kstride_dst[0] = 1; ra.cols = 4096; ra.chunk_len = 4096; ra.chunks = 1; ra.vec = 1;
    Fin fin{FIN_SET, 0.f};
    Fin fadd{FIN_ADD, 0.f};
    prod_old::RowArgs oa; memset(&oa, 0, sizeof(oa));
    oa.nkept = 1; oa.kshape[0] = 4096; oa.kstride_src[0] = 4096; oa.kstride_dst[0] = 1; oa.cols = 4096; oa.chunk_len = 4096; oa.chunks = 1; oa.vec = 1;
    prod_old::Fin ofin{prod_old::FIN_SET, 0.f};
    time_variant("PROD old tile RPC2", st, [&](int i) { prod_old::RowArgs a = oa; a.src = buf[i]; a.dst = part; prod_old::old_tile_kernel<prod_old::MSum, 2><<<2048, 256, 0, st>>>(a, ofin, 4096u); }, bytes);
    time_variant("PROD new tile RPC2 set", st, [&](int i) { RowArgs a = ra; a.src = buf[i]; a.dst = part; new_tile_kernel<MSum, 2><<<2048, 256, 0, st>>>(a, fin, 4096u); }, bytes);
    time_variant("PROD new tile RPC2 accumulate", st, [&](int i) { RowArgs a = ra; a.src = buf[i]; a.dst = part; new_tile_kernel<MSum, 2><<<2048, 256, 0, st>>>(a, fadd, 4096u); }, bytes);
    // sum :axis t as the product launches it: rows = 1, chunks
    RowArgs rt; memset(&rt, 0, sizeof(rt));
    rt.nkept = 0; rt.cols = (int64_t)n; rt.vec = 1; rt.partial = part + 8192; rt.counters = counter;
    time_variant("PROD new rows<256> chunks1024", st, [&](int i) { RowArgs a = rt; a.src = buf[i]; a.dst = out; a.chunks = 1024; a.chunk_len = 16384; new_rows_kernel<MSum, 256><<<1024, 256, 0, st>>>(a, fin); }, bytes);
    time_variant("PROD new rows<512> chunks256", st, [&](int i) { RowArgs a = rt; a.src = buf[i]; a.dst = out; a.chunks = 256; a.chunk_len = 65536; new_rows_kernel<MSum, 512><<<256, 512, 0, st>>>(a, fin); }, bytes);
    time_variant("PROD new rows<512> interleaved 296", st, [&](int i) { RowArgs a = rt; a.src = buf[i]; a.dst = out; a.chunks = 296; a.chunk_len = a.cols; a.interleaved = 1; new_rows_kernel<MSum, 512><<<296, 512, 0, st>>>(a, fin); }, bytes);
    time_variant("PROD new rows<512> interleaved 256", st, [&](int i) { RowArgs a = rt; a.src = buf[i]; a.dst = out; a.chunks = 256; a.chunk_len = a.cols; a.interleaved = 1; new_rows_kernel<MSum, 512><<<256, 512, 0, st>>>(a, fin); }, bytes);
    time_variant("PROD new rows<512> interleaved 148", st, [&](int i) { RowArgs a = rt; a.src = buf[i]; a.dst = out; a.chunks = 148; a.chunk_len = a.cols; a.interleaved = 1; new_rows_kernel<MSum, 512><<<148, 512, 0, st>>>(a, fin); }, bytes);
    time_variant("PROD new rows<256> interleaved 592", st, [&](int i) { RowArgs a = rt; a.src = buf[i]; a.dst = out; a.chunks = 592; a.chunk_len = a.cols; a.interleaved = 1; new_rows_kernel<MSum, 256><<<592, 256, 0, st>>>(a, fin); }, bytes);
  }
  return 0;
}
