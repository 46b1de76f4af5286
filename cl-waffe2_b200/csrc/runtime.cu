// runtime.cu — device selection, storage (the `vec` of a CUDATensor), transfers, timing and
// CUDA-graph capture.  Mirrors the allocator contract of the reference's device classes
// (`source/backends/lisp/tensor.lisp:12-28`, `source/vm/allocation.lisp:71-171`): storage is
// zero-initialised, shared by aliasing, and released through `tensor-finalizer`.
#include <stdarg.h>
#include <stdlib.h>
#include <mutex>
#include <unordered_map>
#include <vector>
#include "common.cuh"

namespace wf2 {

State& S() {
  static State s;
  return s;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("WF2CU_PDL"); on = (e && e[0] == '0') ? 0 : 1; }
  return on != 0;
}

// WF2CU_CARVEOUT=<percent>: every kernel launched through launch_pdl asks for that shared-memory carve-out (experiment: a kernel
// whose carve-out differs from its neighbours' makes the SMs reconfigure at both boundaries)
void carveout_hint(const void* kernel) {
  static int pct = -2;
  if (pct == -2) { const char* e = getenv("WF2CU_CARVEOUT"); pct = e ? atoi(e) : -1; }
  if (pct < 0) return;
  static std::vector<const void*> seen;
  for (const void* k : seen) if (k == kernel) return;
  seen.push_back(kernel);
  cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
}

static thread_local char g_err[1024] = "";

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) in %s at %s:%d", (int)e, cudaGetErrorString(e), what, file, line);
  return WF2CU_ERR_CUDA;
}

static std::vector<void*>& retired_scratch() {
  static std::vector<void*> v;
  return v;
}

int ensure_scratch(size_t bytes) {
  State& s = S();
  bytes = (bytes + 255) & ~size_t(255);
  if (s.scratch_bytes / kBranches >= bytes) return 0;
  // A captured CUDA graph has scratch_base() pointers baked into its reduce / split-K / SCE kernel arguments: the old
  // buffer is retired (kept alive until shutdown), never freed, and growing inside a capture is refused — run the same
  // shapes once before wf2cu_graph_begin (every host path here does: the first step of a CompiledStep is eager).
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (s.stream) cudaStreamIsCapturing(s.stream, &cs);
  if (cs != cudaStreamCaptureStatusNone)
    return set_error(WF2CU_ERR_UNSUPPORTED, "reduction scratch must grow to %zu bytes per branch while a CUDA graph is being captured; "
                     "run the same shapes once before wf2cu_graph_begin", bytes);
  if (s.scratch) retired_scratch().push_back(s.scratch);
  s.scratch = nullptr; s.scratch_bytes = 0;
  size_t want = bytes * kBranches < (size_t(8) << 20) ? (size_t(8) << 20) : bytes * kBranches;
  WF2_CUDA(cudaMalloc(&s.scratch, want));
  s.scratch_bytes = want;
  return 0;
}

void* scratch_base() {
  State& s = S();
  return (char*)s.scratch + (size_t)s.branch * ((s.scratch_bytes / kBranches) & ~size_t(255));
}

unsigned int* counters_base() { return S().counters + (size_t)S().branch * kCountersPerBranch; }

// ---- plan building: drop extent-1 axes, merge neighbours that are jointly contiguous ---------
int build_plan(const wf2cu_view* const* views, int n, Plan* out, const char* who) {
  if (n < 1 || n > kMaxOperands) return set_error(WF2CU_ERR_INVALID, "%s: bad operand count %d", who, n);
  const wf2cu_view* v0 = views[0];
  if (!v0) return set_error(WF2CU_ERR_INVALID, "%s: null view", who);
  int rank = v0->rank;
  if (rank < 0 || rank > kMaxRank) return set_error(WF2CU_ERR_INVALID, "%s: rank %d out of range (max %d)", who, rank, kMaxRank);
  for (int j = 0; j < n; ++j) {
    if (!views[j] || !views[j]->base) return set_error(WF2CU_ERR_INVALID, "%s: operand %d is null", who, j);
    if (views[j]->rank != rank) return set_error(WF2CU_ERR_INVALID, "%s: rank mismatch (%d vs %d)", who, views[j]->rank, rank);
    for (int i = 0; i < rank; ++i) {
      if (views[j]->shape[i] != v0->shape[i])
        return set_error(WF2CU_ERR_INVALID, "%s: shape mismatch on axis %d (%lld vs %lld)", who, i,
                         (long long)views[j]->shape[i], (long long)v0->shape[i]);
      if (v0->shape[i] < 0) return set_error(WF2CU_ERR_INVALID, "%s: negative extent", who);
    }
  }
  Plan p;
  memset(&p, 0, sizeof(p));
  p.nops = n;
  p.numel = 1;
  int r = 0;
  for (int i = 0; i < rank; ++i) {
    int64_t e = v0->shape[i];
    p.numel *= e;
    if (e == 1) continue;
    bool merged = false;
    if (r > 0) {
      bool ok = true;
      for (int j = 0; j < n; ++j)
        if (p.stride[j][r - 1] != views[j]->stride[i] * e) { ok = false; break; }
      if (ok) {
        p.shape[r - 1] *= e;
        for (int j = 0; j < n; ++j) p.stride[j][r - 1] = views[j]->stride[i];
        merged = true;
      }
    }
    if (!merged) {
      p.shape[r] = e;
      for (int j = 0; j < n; ++j) p.stride[j][r] = views[j]->stride[i];
      ++r;
    }
  }
  if (r == 0) {  // scalar-like: one element
    p.shape[0] = 1;
    for (int j = 0; j < n; ++j) p.stride[j][0] = 0;
    r = 1;
  }
  p.rank = r;
  *out = p;
  return 0;
}

// ---- storage registry ---------------------------------------------------------------------
struct Block { size_t bytes; int refs; };
static std::mutex g_mu;
static std::unordered_map<void*, Block>& blocks() {
  static std::unordered_map<void*, Block> m;
  return m;
}

__global__ void touch_kernel(float4* p, size_t n4, float v) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) p[i] = make_float4(v, v, v, v);
}

}  // namespace wf2

using namespace wf2;

extern "C" {

const char* wf2cu_last_error(void) { return g_err; }

int wf2cu_device_count(int* out) {
  if (!out) return set_error(WF2CU_ERR_INVALID, "wf2cu_device_count: null out");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) { *out = 0; return cuda_fail(e, "cudaGetDeviceCount", __FILE__, __LINE__); }
  *out = n;
  return 0;
}

int wf2cu_init(int device) {
  State& s = S();
  if (s.inited && s.device == device) return 0;
  if (s.inited) return set_error(WF2CU_ERR_INVALID, "wf2cu_init: already initialised on device %d", s.device);
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return set_error(WF2CU_ERR_CUDA, "wf2cu_init: no CUDA device visible (%s); this backend has no CPU fallback",
                     e == cudaSuccess ? "count=0" : cudaGetErrorString(e));
  if (device < 0 || device >= n) return set_error(WF2CU_ERR_INVALID, "wf2cu_init: device %d out of range [0,%d)", device, n);
  WF2_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  WF2_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_init: device '%s' is sm_%d%d; this library contains sm_100a code only",
                     prop.name, prop.major, prop.minor);
  s.device = device;
  s.sm_count = prop.multiProcessorCount;
  WF2_CUDA(cudaStreamCreateWithFlags(&s.own_stream, cudaStreamNonBlocking));
  s.stream = s.own_stream;
  WF2_CUDA(cudaEventCreate(&s.ev0));
  WF2_CUDA(cudaEventCreate(&s.ev1));
  WF2_CUDA(cudaMalloc(&s.counters, (size_t)kBranches * kCountersPerBranch * sizeof(unsigned int)));
  WF2_CUDA(cudaMemset(s.counters, 0, (size_t)kBranches * kCountersPerBranch * sizeof(unsigned int)));
  for (int b = 1; b < kBranches; ++b) {
    WF2_CUDA(cudaStreamCreateWithFlags(&s.side[b], cudaStreamNonBlocking));
    WF2_CUDA(cudaEventCreateWithFlags(&s.fork_ev[b], cudaEventDisableTiming));
    WF2_CUDA(cudaEventCreateWithFlags(&s.join_ev[b], cudaEventDisableTiming));
  }
  s.branch = 0;
  int rc = ensure_scratch(size_t(8) << 20);
  if (rc) return rc;
  s.inited = true;
  return 0;
}

int wf2cu_shutdown(void) {
  State& s = S();
  if (!s.inited) return 0;
  cudaStreamSynchronize(s.stream);
  {
    std::lock_guard<std::mutex> lk(g_mu);
    for (auto& kv : blocks()) cudaFree(kv.first);
    blocks().clear();
  }
  gemm_tc_release();
  if (s.scratch) cudaFree(s.scratch);
  for (void* p : retired_scratch()) cudaFree(p);
  retired_scratch().clear();
  if (s.counters) cudaFree(s.counters);
  if (s.l2flush) cudaFree(s.l2flush);
  s.scratch = nullptr; s.scratch_bytes = 0; s.counters = nullptr; s.l2flush = nullptr; s.l2flush_bytes = 0;
  cudaEventDestroy(s.ev0); cudaEventDestroy(s.ev1);
  for (int b = 1; b < kBranches; ++b) {
    if (s.side[b]) { cudaStreamSynchronize(s.side[b]); cudaStreamDestroy(s.side[b]); s.side[b] = nullptr; }
    if (s.fork_ev[b]) { cudaEventDestroy(s.fork_ev[b]); s.fork_ev[b] = nullptr; }
    if (s.join_ev[b]) { cudaEventDestroy(s.join_ev[b]); s.join_ev[b] = nullptr; }
  }
  s.branch = 0;
  cudaStreamDestroy(s.own_stream);
  s.own_stream = s.stream = nullptr;
  s.inited = false;
  s.device = -1;
  s.bytes_in_use = 0;
  return 0;
}

int wf2cu_device_name(char* buf, size_t buflen) {
  WF2_REQUIRE_INIT();
  cudaDeviceProp prop;
  WF2_CUDA(cudaGetDeviceProperties(&prop, S().device));
  snprintf(buf, buflen, "%s (sm_%d%d, %d SMs)", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
  return 0;
}

int wf2cu_sync(void) {
  WF2_REQUIRE_INIT();
  WF2_CUDA(cudaStreamSynchronize(S().stream));
  return 0;
}

int wf2cu_stream(void** out) {
  WF2_REQUIRE_INIT();
  *out = (void*)S().stream;
  return 0;
}

int wf2cu_set_stream(void* st) {
  WF2_REQUIRE_INIT();
  S().stream = st ? (cudaStream_t)st : S().own_stream;
  return 0;
}

int wf2cu_launch_count(int64_t* out) {
  *out = S().launches.load();
  return 0;
}

int wf2cu_malloc(size_t bytes, void** out) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(out != nullptr, "wf2cu_malloc: null out");
  size_t padded = (bytes + 255) & ~size_t(255);   // 16-byte vector access past ragged tails stays in-bounds
  if (padded == 0) padded = 256;
  void* p = nullptr;
  WF2_CUDA(cudaMalloc(&p, padded));
  WF2_CUDA(cudaMemsetAsync(p, 0, padded, S().stream));
  {
    std::lock_guard<std::mutex> lk(g_mu);
    blocks()[p] = Block{padded, 1};
  }
  S().bytes_in_use += padded;
  *out = p;
  return 0;
}

int wf2cu_retain(void* dev) {
  WF2_REQUIRE_INIT();
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = blocks().find(dev);
  if (it == blocks().end()) return set_error(WF2CU_ERR_INVALID, "wf2cu_retain: %p is not a live wf2cu allocation", dev);
  it->second.refs++;
  return 0;
}

int wf2cu_release(void* dev) {
  WF2_REQUIRE_INIT();
  size_t bytes = 0;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = blocks().find(dev);
    if (it == blocks().end()) return set_error(WF2CU_ERR_INVALID, "wf2cu_release: %p is not a live wf2cu allocation (double free?)", dev);
    if (--it->second.refs > 0) return 0;
    bytes = it->second.bytes;
    blocks().erase(it);
  }
  // stream-ordered: work already queued on the buffer finishes first
  WF2_CUDA(cudaStreamSynchronize(S().stream));
  gemm_tc_const_touch(dev, bytes, true);       // planes kept for constant operands in this storage go with it
  WF2_CUDA(cudaFree(dev));
  S().bytes_in_use -= bytes;
  return 0;
}

int wf2cu_free(void* dev) { return wf2cu_release(dev); }

int wf2cu_mem_in_use(size_t* bytes) {
  *bytes = S().bytes_in_use.load();
  return 0;
}

int wf2cu_memset0(void* dev, size_t off, size_t bytes) {
  WF2_REQUIRE_INIT();
  WF2_CUDA(cudaMemsetAsync((char*)dev + off, 0, bytes, S().stream));
  gemm_tc_const_touch((char*)dev + off, bytes, false);
  return 0;
}

int wf2cu_h2d(void* dev, size_t off, const void* host, size_t bytes) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dev && (host || bytes == 0), "wf2cu_h2d: null pointer");
  if (bytes == 0) return 0;
  WF2_CUDA(cudaMemcpyAsync((char*)dev + off, host, bytes, cudaMemcpyHostToDevice, S().stream));
  gemm_tc_const_touch((char*)dev + off, bytes, false);   // a new value for a constant operand: its planes are split again
  return 0;
}

int wf2cu_d2h(const void* dev, size_t off, void* host, size_t bytes) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dev && (host || bytes == 0), "wf2cu_d2h: null pointer");
  if (bytes) WF2_CUDA(cudaMemcpyAsync(host, (const char*)dev + off, bytes, cudaMemcpyDeviceToHost, S().stream));
  WF2_CUDA(cudaStreamSynchronize(S().stream));
  return 0;
}

int wf2cu_d2h_async(const void* dev, size_t off, void* host, size_t bytes) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dev && (host || bytes == 0), "wf2cu_d2h_async: null pointer");
  if (bytes) WF2_CUDA(cudaMemcpyAsync(host, (const char*)dev + off, bytes, cudaMemcpyDeviceToHost, S().stream));
  return 0;
}

int wf2cu_d2d(void* dst, size_t doff, const void* src, size_t soff, size_t bytes) {
  WF2_REQUIRE_INIT();
  if (bytes) WF2_CUDA(cudaMemcpyAsync((char*)dst + doff, (const char*)src + soff, bytes, cudaMemcpyDeviceToDevice, S().stream));
  if (bytes) gemm_tc_const_touch((char*)dst + doff, bytes, false);
  return 0;
}

int wf2cu_get_f32(const void* dev, int64_t index, float* out) {
  return wf2cu_d2h(dev, (size_t)index * 4, out, 4);
}

int wf2cu_set_f32(void* dev, int64_t index, float value) {
  WF2_REQUIRE_INIT();
  // value is copied by the runtime before returning for pageable sources
  WF2_CUDA(cudaMemcpyAsync((char*)dev + index * 4, &value, 4, cudaMemcpyHostToDevice, S().stream));
  gemm_tc_const_touch((char*)dev + index * 4, 4, false);   // (setf vref) into a constant operand / a tensor with producer-written planes
  return 0;
}

int wf2cu_host_alloc(size_t bytes, void** out) {
  WF2_REQUIRE_INIT();
  WF2_CUDA(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
  return 0;
}

int wf2cu_host_free(void* host) {
  WF2_REQUIRE_INIT();
  WF2_CUDA(cudaFreeHost(host));
  return 0;
}

int wf2cu_timer_start(void) {
  WF2_REQUIRE_INIT();
  WF2_CUDA(cudaEventRecord(S().ev0, S().stream));
  return 0;
}

int wf2cu_timer_stop(float* ms) {
  WF2_REQUIRE_INIT();
  WF2_CUDA(cudaEventRecord(S().ev1, S().stream));
  WF2_CUDA(cudaEventSynchronize(S().ev1));
  WF2_CUDA(cudaEventElapsedTime(ms, S().ev0, S().ev1));
  return 0;
}

int wf2cu_l2_flush(void) {
  WF2_REQUIRE_INIT();
  State& s = S();
  if (!s.l2flush) {
    s.l2flush_bytes = size_t(256) << 20;  // 2x the 126 MB L2
    WF2_CUDA(cudaMalloc(&s.l2flush, s.l2flush_bytes));
  }
  touch_kernel<<<s.sm_count * 8, 256, 0, s.stream>>>((float4*)s.l2flush, s.l2flush_bytes / 16, 1.0f);
  // not counted as a product launch: it is measurement hygiene
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "l2_flush", __FILE__, __LINE__);
  return 0;
}

// launches recorded while capturing are credited on every replay
static std::unordered_map<void*, int64_t>& graph_launches() {
  static std::unordered_map<void*, int64_t> m;
  return m;
}
static int64_t g_capture_start = 0;

int wf2cu_graph_begin(void) {
  WF2_REQUIRE_INIT();
  WF2_CUDA(cudaStreamBeginCapture(S().stream, cudaStreamCaptureModeThreadLocal));
  g_capture_start = S().launches.load();
  return 0;
}

int wf2cu_graph_end(void** out_exec) {
  WF2_REQUIRE_INIT();
  cudaGraph_t g = nullptr;
  WF2_CUDA(cudaStreamEndCapture(S().stream, &g));
  cudaGraphExec_t ex = nullptr;
  cudaError_t e = cudaGraphInstantiate(&ex, g, 0);
  cudaGraphDestroy(g);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGraphInstantiate", __FILE__, __LINE__);
  const int64_t captured = S().launches.load() - g_capture_start;
  S().launches.fetch_sub(captured);  // nothing ran yet
  {
    std::lock_guard<std::mutex> lk(g_mu);
    graph_launches()[(void*)ex] = captured;
  }
  *out_exec = (void*)ex;
  return 0;
}

int wf2cu_graph_launch(void* exec) {
  WF2_REQUIRE_INIT();
  WF2_CUDA(cudaGraphLaunch((cudaGraphExec_t)exec, S().stream));
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = graph_launches().find(exec);
    if (it != graph_launches().end()) S().launches.fetch_add(it->second);
  }
  return 0;
}

int wf2cu_graph_destroy(void* exec) {
  WF2_REQUIRE_INIT();
  WF2_CUDA(cudaGraphExecDestroy((cudaGraphExec_t)exec));
  {
    std::lock_guard<std::mutex> lk(g_mu);
    graph_launches().erase(exec);
  }
  return 0;
}

// ---- side branches: independent launches of a compiled step on their own stream ------------------
// The reference's VM executes a compiled instruction sequence strictly in order (vm/vm.lisp:260-289).  On a
// GPU, instructions with no data dependence between them (the weight-gradient GEMM and the bias-gradient
// reduction of a layer vs. the input-gradient GEMM the next backward node waits for) can overlap: each is
// far too small to fill 148 SMs.  begin(b): launches from now on go to side stream b, ordered after
// everything issued so far; end(b): back to the main stream; join(b): the main stream waits for branch b.
// Inside wf2cu_graph_begin/end these become graph edges (fork/join), so the replayed graph runs the branches
// concurrently.  Each branch has its own quarter of the reduction scratch and tickets.
int wf2cu_branch_begin(int b) {
  WF2_REQUIRE_INIT();
  State& s = S();
  WF2_CHECK(b >= 1 && b < kBranches, "wf2cu_branch_begin: branch %d not in 1..%d", b, kBranches - 1);
  WF2_CHECK(s.branch == 0, "wf2cu_branch_begin: branch %d is still open (branches do not nest)", s.branch);
  s.main_stream = s.stream;
  WF2_CUDA(cudaEventRecord(s.fork_ev[b], s.main_stream));
  WF2_CUDA(cudaStreamWaitEvent(s.side[b], s.fork_ev[b], 0));
  s.stream = s.side[b];
  s.branch = b;
  return 0;
}

int wf2cu_branch_end(int b) {
  WF2_REQUIRE_INIT();
  State& s = S();
  WF2_CHECK(b >= 1 && b < kBranches && s.branch == b, "wf2cu_branch_end: branch %d is not the open branch (%d)", b, s.branch);
  WF2_CUDA(cudaEventRecord(s.join_ev[b], s.side[b]));
  s.stream = s.main_stream;
  s.branch = 0;
  return 0;
}

int wf2cu_branch_join(int b) {
  WF2_REQUIRE_INIT();
  State& s = S();
  WF2_CHECK(b >= 1 && b < kBranches, "wf2cu_branch_join: branch %d not in 1..%d", b, kBranches - 1);
  WF2_CHECK(s.branch == 0, "wf2cu_branch_join: close branch %d first", s.branch);
  WF2_CUDA(cudaStreamWaitEvent(s.stream, s.join_ev[b], 0));
  return 0;
}

}  // extern "C"
