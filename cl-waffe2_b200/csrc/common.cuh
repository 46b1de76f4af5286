// common.cuh — shared state, error plumbing and launch helpers for libwaffe2cuda.so
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <atomic>
#include "../../include/waffe2cuda.h"

namespace wf2 {

constexpr int kMaxRank = WF2CU_MAX_RANK;
constexpr int kNumSMs = 148;  // B200
constexpr int kBranches = 4;   // main + 3 side streams; each owns a quarter of the scratch buffer and of the tickets
constexpr int kCountersPerBranch = 4096;

struct State {
  bool inited = false;
  int device = -1;
  int sm_count = kNumSMs;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  void* scratch = nullptr;       // reduction partials
  size_t scratch_bytes = 0;
  unsigned int* counters = nullptr;  // last-block tickets (zeroed, self-resetting)
  void* l2flush = nullptr;
  size_t l2flush_bytes = 0;
  std::atomic<int64_t> launches{0};
  std::atomic<size_t> bytes_in_use{0};
  int gemm_mode = 0;
  int gemm_last_path = 0;        // WF2CU_GEMM_PATH_* of the last wf2cu_gemm / wf2cu_gemm_ex call (wf2cu_gemm_last_path)
  // side branches (wf2cu_branch_begin/end/join): independent work of a compiled step on its own stream
  cudaStream_t main_stream = nullptr;             // where launches return to after wf2cu_branch_end
  cudaStream_t side[kBranches] = {};
  cudaEvent_t fork_ev[kBranches] = {}, join_ev[kBranches] = {};
  int branch = 0;                                 // 0 = main, 1..kBranches-1 = side
  // operand planes written by a PRODUCER for the MatmulNode that follows it (wf2cu_emit_operand_planes; gemm_tc.cu):
  // valid from the producer's launch until the next launch of any kind (post_launch clears it; the consuming product looks first)
  struct FreshPlanes { const float* ptr = nullptr; int64_t rows = 0, k = 0, ld = 0; float* lo = nullptr; } fresh;
  int emit_next = 0;                              // one-shot: the next producer call also writes its result's planes
};
State& S();

int set_error(int code, const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define WF2_CUDA(call)                                                     \
  do {                                                                     \
    cudaError_t _e = (call);                                               \
    if (_e != cudaSuccess) return wf2::cuda_fail(_e, #call, __FILE__, __LINE__); \
  } while (0)

#define WF2_REQUIRE_INIT()                                                  \
  do {                                                                     \
    if (!wf2::S().inited)                                                  \
      return wf2::set_error(WF2CU_ERR_NOT_INITIALISED,                     \
                            "libwaffe2cuda: wf2cu_init has not succeeded (no B200 / no CPU fallback)"); \
  } while (0)

#define WF2_CHECK(cond, ...)                                               \
  do {                                                                     \
    if (!(cond)) return wf2::set_error(WF2CU_ERR_INVALID, __VA_ARGS__);    \
  } while (0)

// call after every kernel launch
inline int post_launch(const char* name) {
  S().launches.fetch_add(1, std::memory_order_relaxed);
  S().fresh.ptr = nullptr;                         // whatever was launched may have written the tensor the planes belong to
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, name, __FILE__, __LINE__);
  return 0;
}

int ensure_scratch(size_t bytes);           // per branch: the buffer holds kBranches regions of at least `bytes`
void* scratch_base();                        // the current branch's region (valid after ensure_scratch)
unsigned int* counters_base();               // the current branch's kCountersPerBranch tickets
void gemm_tc_release();   // gemm_tc.cu: frees the hi/lo plane workspaces
// gemm_tc.cu: storage [ptr, ptr + bytes) was overwritten by a transfer (released = false: planes kept for a constant operand
// there are split again on the next product) or freed (released = true: planes and the constant declaration are dropped)
void gemm_tc_const_touch(const void* ptr, size_t bytes, bool released);
// gemm_tc.cu: a producer that was asked to (S().emit_next, cleared here) gets the buffer for the mixed bf16 operand plane of its
// (rows x k) fp32 result — layout of split_planes_direct_kernel with k-blocks of 16 — or nullptr when the request cannot be
// met (k not a multiple of 32, allocation needed inside a graph capture, side branch).  After its launch it publishes them.
float* planes_emit_begin(int64_t rows, int64_t k);
void planes_emit_publish(const float* result, int64_t rows, int64_t k, int64_t ld, float* lo);
#ifdef __CUDACC__
// the two 16-bit halves a producer stores for element x: bf16_rn(x) and bf16_rn(x - trunc_tf32(x)); 0 / 0 for inf and nan
// (exactly split_bf<true> of gemm_tc.cu, so a produced plane equals the pre-pass's bit for bit)
__device__ __forceinline__ void plane_bits(float x, uint32_t& hb, uint32_t& lb) {
  const float hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  if (fabsf(hi) <= 3.0e38f) {
    asm("{\n\t.reg .b16 t;\n\tcvt.rn.bf16.f32 t, %1;\n\tcvt.u32.u16 %0, t;\n\t}" : "=r"(hb) : "f"(x));
    asm("{\n\t.reg .b16 t;\n\tcvt.rn.bf16.f32 t, %1;\n\tcvt.u32.u16 %0, t;\n\t}" : "=r"(lb) : "f"(__fsub_rn(x, hi)));
  } else { hb = 0; lb = 0; }
}
// four consecutive elements (columns c .. c + 3, c a multiple of 4) of a row of the plane: two 8-byte stores
__device__ __forceinline__ void plane_store4(float* plane_row, int c, float x0, float x1, float x2, float x3) {
  uint32_t h0, h1, h2, h3, l0, l1, l2, l3;
  plane_bits(x0, h0, l0); plane_bits(x1, h1, l1); plane_bits(x2, h2, l2); plane_bits(x3, h3, l3);
  uint16_t* dst = reinterpret_cast<uint16_t*>(plane_row) + (c >> 4) * 32 + (c & 15);
  *reinterpret_cast<uint2*>(dst) = make_uint2(h0 | (h1 << 16), h2 | (h3 << 16));
  *reinterpret_cast<uint2*>(dst + 16) = make_uint2(l0 | (l1 << 16), l2 | (l3 << 16));
}
#endif

// ---- programmatic dependent launch (PDL) -------------------------------------------------------------
// Back-to-back kernels of a compiled instruction sequence are serialised by the stream (or by the edges of the captured
// graph): kernel N+1 is launched only when the last CTA of kernel N has retired, so every boundary costs a launch
// latency plus the ramp of the new grid — 1-2 us against 12-30 us kernels.  A kernel launched through launch_pdl may
// start its CTAs while its predecessor drains; it calls pdl_prologue() before its first global memory access, which
// (1) lets ITS successor do the same and (2) blocks until the predecessor grid has completed and its writes are
// visible.  Without the launch attribute both instructions are no-ops, so the same kernel works either way.
// WF2CU_PDL=0 turns the attribute off (A/B measurements).
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_prologue() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
#endif
bool pdl_enabled();
void carveout_hint(const void* kernel);
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  carveout_hint(reinterpret_cast<const void*>(kernel));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

inline float* fptr(const wf2cu_view* v) { return reinterpret_cast<float*>(v->base) + v->offset; }

// ---- iteration-space plan shared by all strided kernels -------------------------------
// After normalisation: size-1 axes dropped, adjacent axes merged when every operand allows it.
constexpr int kMaxOperands = 8;
struct Plan {
  int rank;                       // >= 1
  int nops;
  int64_t shape[kMaxRank];
  int64_t stride[kMaxOperands][kMaxRank];
  int64_t numel;
};

// Build a merged plan over n equal-shape views. Returns 0 or error.
int build_plan(const wf2cu_view* const* views, int n, Plan* out, const char* who);

inline int64_t view_numel(const wf2cu_view* v) {
  int64_t n = 1;
  for (int i = 0; i < v->rank; ++i) n *= v->shape[i];
  return n;
}

}  // namespace wf2
