// p2p.cu — the data-parallel step's exchange + optimizer as ONE kernel over NVLink peer memory.
//
// after backward every rank holds the gradient bucket of its batch shard; the global step
// is allreduce(sum) x 1/world, then the identical Adam update on every rank.  The NCCL route
// (comm.cu) is a collective launch followed by the optimizer launch.  Here every rank maps its peers'
// gradient arenas (CUDA IPC, one process per GPU; NVSwitch gives each pair full bandwidth), and one
// kernel per rank
//   1. tells its peers "my gradients are complete" and waits for theirs  (st.release.sys / ld.acquire.sys
//      on a small signal page in each rank's memory),
//   2. owns shard `rank` of the bucket (1/world of it): reads that shard from ALL ranks' arenas with
//      128-bit loads over NVLink, sums in rank order, scales by 1/world, applies Adam with its own
//      m/v shard (optimizer state and optimizer math are sharded world ways), and stores the updated
//      parameters into EVERY replica's parameter arena (remote 128-bit stores) — reduce-scatter,
//      optimizer and all-gather in one pass, 2 x 1/world of the bucket per rank over NVLink
//      (0.27 MB at N = 8 instead of the 7.5 MB a read-everything scheme moves),
//   3. tells its peers "my shard is written, I am done reading" and waits for theirs, so the next
//      forward sees the new parameters and the next backward may overwrite the gradient arena.
// Each element is computed by exactly one rank and copied, so the replicas stay bit-identical.
// The reference has no distributed code; this file is new work.
//
// Two kernels implement it.  `allreduce_adam_kernel` (PULL, any bucket size) is the description above: two flag exchanges
// (release / acquire at system scope: each one costs the fence's round trip on top of the flag's flight) around one round
// trip of loads and one of stores — measured ~20 us on top of a 90 us step at N = 2..8 (r02_scale_n*.json), all of it latency.
// `allreduce_adam_ll_kernel` (PUSH, buckets up to kLLMaxElems floats) removes every fence and every flag from the path by
// sending the FLAG WITH THE DATA, the way NCCL's LL protocol does: each 16-byte store carries {value, epoch, value, epoch};
// an 8-byte half arrives whole, so a receiver that sees the current epoch in both halves holds both values.  Per step:
//   A. every rank pushes each gradient element to the staging area of the rank that owns it   (posted stores, one flight)
//   B. the owner polls its staging area (local memory), sums in rank order, x 1/world, Adam, writes its own replica and
//      pushes the new parameters to every peer's parameter staging area                       (posted stores, one flight)
//   C. every rank polls its parameter staging area (local memory) and copies into its replica.
// Two one-way flights instead of two fenced flag exchanges plus two round trips.  No "done" flags are needed: a peer can only
// overwrite a staging slot of step k with data of step k + 1 after it has finished step k, which needed this rank's parameters
// of step k, which this rank sent after it had read that slot (the argument is spelled out in DESIGN.md §3.6).  Same
// summation order, same Adam arithmetic: results are bit-identical to the pull kernel and to the numpy emulation.
#include "common.cuh"

namespace wf2 {

constexpr int kMaxPeers = 8;
constexpr unsigned long long kSpinLimitNs = 10000000000ull;  // 10 s: a lost peer must not hang the GPU, ordinary host skew must not trip it

// A timeout is STICKY and FATAL for the attachment: the rank that gives up sets its own flag (device word local[2] and a
// mapped pinned host word the host polls without a sync) and writes `abort` into every peer's signal page; a rank that
// sees `abort` (while spinning or at kernel entry) does the same.  From then on every rank's kernel returns at entry without
// touching flags, gradients or parameters, so no replica is updated from a half-arrived exchange and no peer reads an
// arena its owner has moved on from; the host raises on the next step (nn.CompiledStep.step polls wf2cu_p2p_status).
struct SignalPage {                 // lives in each rank's device memory, written by its peers
  uint32_t ready[16];               // ready[r] = epoch for which rank r's gradients are complete
  uint32_t done[16];                // done[r]  = epoch for which rank r has finished reading my arena
  uint32_t abort;                   // != 0: some rank timed out; the attachment is dead
};
// The exported window: the signal page, then the two staging areas of the LL kernel (8 bytes per fp32 element each):
//   gstage[r][j]  gradient float4 j of MY shard as sent by rank r   (world x per x 32 bytes)
//   pstage[i]     parameter float4 i of the bucket as sent by its owner (n4 x 32 bytes)
constexpr size_t kSignalBytes = 4096;
constexpr int64_t kLLMaxElems = 1 << 20;                                   // fp32 elements per bucket on the LL path
constexpr size_t kLLStageBytes = (size_t)(kLLMaxElems + 2048) * 8;         // one staging area (padding: every sender's row is rounded up to 32 float4)
constexpr size_t kWindowBytes = kSignalBytes + 2 * kLLStageBytes;
constexpr int kLLMaxCtas = 1024;
constexpr size_t kLocalBytes = 256 + kLLMaxCtas * sizeof(uint32_t);

struct P2PState {
  int rank = -1, world = 0;
  void* peer_grad_base[kMaxPeers] = {};     // mapped allocation bases (self: the local pointer)
  void* peer_param_base[kMaxPeers] = {};
  bool opened_param[kMaxPeers] = {};
  SignalPage* peer_signal[kMaxPeers] = {};
  bool opened_grad[kMaxPeers] = {}, opened_signal[kMaxPeers] = {};
  SignalPage* my_signal = nullptr;
  uint32_t* local = nullptr;                // [0] epoch, [1] ticket, [2] timeout flag (sticky); [64 ..) the LL kernel's per-CTA epochs
  uint32_t* host_flag = nullptr;            // mapped pinned host word, written by the kernel on a timeout (host polls it, no sync)
  uint32_t* host_flag_dev = nullptr;        // its device address
  int64_t ll_grid = 0;                      // grid of the LL kernel in this attachment (its per-CTA epochs need a fixed grid)
};
static P2PState g_p2p;

struct P2PArgs {
  float* param; float* m; float* v;
  const float* lr_dev;
  float lr_t, b1, b2, eps, gscale;
  int64_t n;                                // elements (the arenas are padded to a multiple of 4)
  int world, rank;
  const float* grads[kMaxPeers];
  float* params[kMaxPeers];                 // every replica's parameter arena (params[rank] == param)
  SignalPage* signals[kMaxPeers];
  uint32_t* local;
  uint32_t* host_flag;
  const float* grad;                        // this rank's gradient arena (LL kernel)
  uint32_t* ll_epochs;                      // LL kernel: one epoch counter per CTA (device memory of this rank)
  unsigned long long* trace;                // WF2CU_P2P_TRACE=1: phase stamps of the LL kernel (see p2p_trace_report), else null
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// wait until *p has reached epoch e (wrap-safe); false on timeout or when a peer has declared the attachment dead
__device__ __forceinline__ bool wait_epoch(const uint32_t* p, uint32_t e, const uint32_t* abort_word) {
  const unsigned long long t0 = now_ns();
  uint32_t spins = 0;
  while ((int32_t)(ld_acquire_sys(p) - e) < 0) {
    if ((++spins & 63u) == 0 && ld_acquire_sys(abort_word) != 0) return false;
    if (now_ns() - t0 > kSpinLimitNs) return false;
    __nanosleep(64);
  }
  return true;
}
// this rank gives up: sticky local flag, the host's word, and `abort` in every peer's page (and its own)
__device__ __forceinline__ void declare_dead(const P2PArgs& a) {
  a.local[2] = 1;
  st_release_sys(a.host_flag, 1u);
  for (int r = 0; r < a.world; ++r) st_release_sys(&a.signals[r]->abort, 1u);
}
__device__ __forceinline__ float4 ld_peer(const float4* p) {
  float4 v;   // L1 no-allocate: every line is read once, and a stale line must never be served next step
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

__device__ __forceinline__ float adam_one(float gi, float& m, float& v, float p, float lr_t, float omb1, float omb2, float eps) {
  // exactly adam_kernel (misc.cu), i.e. optimizers/impls/adam.lisp:45-67
  const float mi = __fadd_rn(m, __fmul_rn(__fsub_rn(gi, m), omb1));
  const float vi = __fadd_rn(v, __fmul_rn(omb2, __fsub_rn(__fmul_rn(gi, gi), v)));
  m = mi; v = vi;
  return __fsub_rn(p, __fdiv_rn(__fmul_rn(mi, lr_t), __fadd_rn(eps, __fsqrt_rn(vi))));
}

__global__ void __launch_bounds__(256) allreduce_adam_kernel(const __grid_constant__ P2PArgs a) {
  __shared__ uint32_t epoch_s;
  __shared__ int alive_s, ready_s;
  __shared__ bool last_s;
  if (threadIdx.x == 0) {
    epoch_s = a.local[0] + 1;
    // a dead attachment stays dead: no flags, no loads from peers, no parameter stores
    alive_s = (a.local[2] == 0 && ld_acquire_sys(&a.signals[a.rank]->abort) == 0) ? 1 : 0;
    ready_s = 1;
  }
  __syncthreads();
  if (!alive_s) {
    if (blockIdx.x == 0 && threadIdx.x == 0) declare_dead(a);
    return;
  }
  const uint32_t e = epoch_s;
  // ---- 1. gradients complete everywhere?  (the kernel follows this rank's backward in stream order)
  if (blockIdx.x == 0 && threadIdx.x < a.world) st_release_sys(&a.signals[threadIdx.x]->ready[a.rank], e);
  if (threadIdx.x < a.world)
    if (!wait_epoch(&a.signals[a.rank]->ready[threadIdx.x], e, &a.signals[a.rank]->abort)) { ready_s = 0; declare_dead(a); }
  __syncthreads();
  if (!ready_s) return;              // no update, no ticket, no done flag, the epoch does not advance: every later launch returns at entry
  // ---- 2. my shard: sum over ranks in rank order, x 1/world, Adam, parameters out to every replica
  {
    const float lr_t = a.lr_dev ? *a.lr_dev : a.lr_t;
    const float omb1 = __fsub_rn(1.0f, a.b1), omb2 = __fsub_rn(1.0f, a.b2);
    const int64_t n4 = (a.n + 3) >> 2;
    const int64_t per = (n4 + a.world - 1) / a.world;
    const int64_t lo = per * a.rank, hi = lo + per < n4 ? lo + per : n4;
    const int64_t stride = (int64_t)gridDim.x * 256;
    for (int64_t base = lo + (int64_t)blockIdx.x * 256; base < hi; base += stride) {   // uniform trip count per CTA
      const int64_t i = base + threadIdx.x;
      const bool live = i < hi;
      float4 x[kMaxPeers];
#pragma unroll
      for (int r = 0; r < kMaxPeers; ++r)
        if (r < a.world && live) x[r] = ld_peer(reinterpret_cast<const float4*>(a.grads[r]) + i);
      // All remote loads are issued before the barrier and consumed after it, so the (up to 8) NVLink
      // round trips of a thread overlap instead of being serialised by the register allocator.
      __syncthreads();
      if (live) {
        float4 g = x[0];
#pragma unroll
        for (int r = 1; r < kMaxPeers; ++r)
          if (r < a.world) { g.x = __fadd_rn(g.x, x[r].x); g.y = __fadd_rn(g.y, x[r].y); g.z = __fadd_rn(g.z, x[r].z); g.w = __fadd_rn(g.w, x[r].w); }
        if (a.gscale != 1.0f) { g.x = __fmul_rn(g.x, a.gscale); g.y = __fmul_rn(g.y, a.gscale); g.z = __fmul_rn(g.z, a.gscale); g.w = __fmul_rn(g.w, a.gscale); }
        float4 m = reinterpret_cast<float4*>(a.m)[i], v = reinterpret_cast<float4*>(a.v)[i], p = reinterpret_cast<float4*>(a.param)[i];
        p.x = adam_one(g.x, m.x, v.x, p.x, lr_t, omb1, omb2, a.eps);
        p.y = adam_one(g.y, m.y, v.y, p.y, lr_t, omb1, omb2, a.eps);
        p.z = adam_one(g.z, m.z, v.z, p.z, lr_t, omb1, omb2, a.eps);
        p.w = adam_one(g.w, m.w, v.w, p.w, lr_t, omb1, omb2, a.eps);
        reinterpret_cast<float4*>(a.m)[i] = m; reinterpret_cast<float4*>(a.v)[i] = v;
#pragma unroll
        for (int r = 0; r < kMaxPeers; ++r)
          if (r < a.world) reinterpret_cast<float4*>(a.params[r])[i] = p;
      }
    }
  }
  // ---- 3. last CTA of this rank: "shard written, done reading" to everyone, wait for everyone, advance the epoch
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();          // this CTA's remote parameter stores are ordered before its ticket
    const unsigned int t = atomicAdd(&a.local[1], 1u);
    last_s = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!last_s) return;
  __threadfence_system();            // ... and all CTAs' stores before the done flags below
  if (threadIdx.x < a.world) st_release_sys(&a.signals[threadIdx.x]->done[a.rank], e);
  __shared__ int done_ok_s;
  if (threadIdx.x == 0) done_ok_s = 1;
  __syncthreads();
  if (threadIdx.x < a.world)
    if (!wait_epoch(&a.signals[a.rank]->done[threadIdx.x], e, &a.signals[a.rank]->abort)) { done_ok_s = 0; declare_dead(a); }
  __syncthreads();
  if (threadIdx.x == 0) { a.local[1] = 0; if (done_ok_s) a.local[0] = e; }
}


// ---------------------------------------------------------------------------------------------------- LL (push) kernel
constexpr int kLLWord1 = 32;   // distance between the two LL words of a float4, in uint4 units (see ll_slot)
__device__ __forceinline__ void st_ll(uint4* p, float a, float b, uint32_t e) {       // {a, e, b, e}: each 8-byte half arrives whole
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(__float_as_uint(a)), "r"(e), "r"(__float_as_uint(b)), "r"(e) : "memory");
}
__device__ __forceinline__ uint4 ld_ll(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
// one float4 = two LL words; false on timeout / abort
__device__ __forceinline__ bool wait_ll(const uint4* p, uint32_t e, float4& out, const uint32_t* abort_word) {
  uint4 a = ld_ll(p), b = ld_ll(p + kLLWord1);
  if (a.y != e || a.w != e || b.y != e || b.w != e) {
    const unsigned long long t0 = now_ns();
    uint32_t spins = 0;
    for (;;) {
      a = ld_ll(p); b = ld_ll(p + kLLWord1);
      if (a.y == e && a.w == e && b.y == e && b.w == e) break;
      if ((++spins & 255u) == 0) {
        if (ld_acquire_sys(abort_word) != 0 || now_ns() - t0 > kSpinLimitNs) return false;
      }
      if (spins > 64) __nanosleep(32);
    }
  }
  out = make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x), __uint_as_float(b.z));
  return true;
}
// Staging layout: float4 number q of an area = two LL words; the words of 32 consecutive float4 are laid out as 512 bytes of
// first words followed by 512 bytes of second words, so that each of a warp's two store instructions writes four WHOLE 128-byte
// lines (a 16-byte store per lane at a 32-byte pitch would leave every line of both instructions half written: twice the
// NVLink packets for the same bytes).  kLLWord1 = distance between the two words in uint4 units.
__device__ __forceinline__ int64_t ll_slot(int64_t q) { return (q >> 5) * 64 + (q & 31); }
__device__ __forceinline__ uint4* ll_gstage(SignalPage* win, int from_rank, int64_t per32, int64_t j) {
  return reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(win) + kSignalBytes) + ll_slot((int64_t)from_rank * per32 + j);
}
__device__ __forceinline__ uint4* ll_pstage(SignalPage* win, int64_t i) {
  return reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(win) + kSignalBytes + kLLStageBytes) + ll_slot(i);
}

// The grid must be co-resident (every CTA pushes before it polls, but a CTA that is never scheduled never pushes): the host
// launches at most 2 CTAs of 256 threads per SM and the loops stride over the grid.
// What the kernel costs with nothing on the wire decides how much of a 90 us step it adds (scripts/p2p_world1.py, one rank
// attached to itself: plain Adam launch 90.9 us per step, the pull kernel 107.0, the first LL kernel 96.6), so the entry and the
// exit are as short as Adam's: every CTA keeps its OWN epoch counter (read by all its threads together with the operands of
// its first element — no leading thread, no barrier, no ticket at the end), the abort word of the signal page is looked at only
// while polling, and the launch is a programmatic dependent one.
__global__ void __launch_bounds__(256) allreduce_adam_ll_kernel(const __grid_constant__ P2PArgs a) {
  pdl_prologue();
  const uint32_t e = a.ll_epochs[blockIdx.x] + 1;           // the same value in every CTA of every rank: one increment per launch
  const uint32_t dead = a.local[2];
  const int64_t n4 = (a.n + 3) >> 2;
  const int64_t per = (n4 + a.world - 1) / a.world;
  const int64_t lo = per * a.rank, hi = lo + per < n4 ? lo + per : n4;
  const int64_t per32 = (per + 31) & ~int64_t(31);          // pitch of one sender's row in a gradient staging area
  const int64_t tid = (int64_t)blockIdx.x * 256 + threadIdx.x, nthr = (int64_t)gridDim.x * 256;
  // operands of this thread's first element of B: in flight while A pushes
  float4 m0 = make_float4(0.f, 0.f, 0.f, 0.f), v0 = m0, p0 = m0, own0 = m0;
  if (lo + tid < hi) {
    const int64_t i = lo + tid;
    m0 = reinterpret_cast<const float4*>(a.m)[i]; v0 = reinterpret_cast<const float4*>(a.v)[i];
    p0 = reinterpret_cast<const float4*>(a.param)[i]; own0 = reinterpret_cast<const float4*>(a.grad)[i];
  }
  const float lr_t = a.lr_dev ? *a.lr_dev : a.lr_t;
  if (dead) {                                               // a dead attachment stays dead: no stores, no polling
    if (tid == 0) declare_dead(a);
    return;
  }
  if (a.trace && threadIdx.x == 0) atomicMin(&a.trace[0], now_ns());
  SignalPage* me = a.signals[a.rank];
  const uint32_t* abort_word = &me->abort;
  bool ok = true;
  // ---- A. my gradients go to their owners (posted stores over NVLink; my own shard is read in place in B)
  for (int64_t i = tid; i < n4; i += nthr) {
    const int o = (int)(i / per);
    if (o == a.rank) continue;
    const float4 g = reinterpret_cast<const float4*>(a.grad)[i];
    uint4* dst = ll_gstage(a.signals[o], a.rank, per32, i - (int64_t)o * per);
    st_ll(dst, g.x, g.y, e);
    st_ll(dst + kLLWord1, g.z, g.w, e);
  }
  if (a.trace) { __syncthreads(); if (threadIdx.x == 0) atomicMax(&a.trace[1], now_ns()); }
  // ---- B. my shard: sum over ranks in rank order, x 1/world, Adam, my replica, and out to every peer
  {
    const float omb1 = __fsub_rn(1.0f, a.b1), omb2 = __fsub_rn(1.0f, a.b2);
    bool first = true;
    for (int64_t i = lo + tid; i < hi; i += nthr, first = false) {
      const int64_t j = i - lo;
      float4 m, v, p, own;
      if (first) { m = m0; v = v0; p = p0; own = own0; }
      else {
        m = reinterpret_cast<const float4*>(a.m)[i]; v = reinterpret_cast<const float4*>(a.v)[i];
        p = reinterpret_cast<const float4*>(a.param)[i]; own = reinterpret_cast<const float4*>(a.grad)[i];
      }
      // All peers' words are requested at once and the flags checked afterwards: polling peer after peer made phase B a chain of
      // world - 1 dependent L2 round trips although the gradients arrive at about the same time.
      float4 x[kMaxPeers];
      uint32_t pending = 0;
#pragma unroll
      for (int r = 0; r < kMaxPeers; ++r)
        if (r < a.world && r != a.rank) pending |= 1u << r;
      {
        const unsigned long long t0 = now_ns();
        uint32_t spins = 0;
        while (pending) {
          uint4 wa[kMaxPeers], wb[kMaxPeers];
#pragma unroll
          for (int r = 0; r < kMaxPeers; ++r)
            if (pending & (1u << r)) { const uint4* q = ll_gstage(me, r, per32, j); wa[r] = ld_ll(q); wb[r] = ld_ll(q + kLLWord1); }
#pragma unroll
          for (int r = 0; r < kMaxPeers; ++r)
            if ((pending & (1u << r)) && wa[r].y == e && wa[r].w == e && wb[r].y == e && wb[r].w == e) {
              x[r] = make_float4(__uint_as_float(wa[r].x), __uint_as_float(wa[r].z), __uint_as_float(wb[r].x), __uint_as_float(wb[r].z));
              pending &= ~(1u << r);
            }
          if (pending && (++spins & 255u) == 0 && (ld_acquire_sys(abort_word) != 0 || now_ns() - t0 > kSpinLimitNs)) { ok = false; break; }
          if (pending && spins > 64) __nanosleep(32);
        }
      }
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < kMaxPeers; ++r)
        if (r < a.world) {
          const float4 xr = r == a.rank ? own : x[r];
          if (r == 0) g = xr;
          else { g.x = __fadd_rn(g.x, xr.x); g.y = __fadd_rn(g.y, xr.y); g.z = __fadd_rn(g.z, xr.z); g.w = __fadd_rn(g.w, xr.w); }
        }
      if (!ok) break;
      if (a.gscale != 1.0f) { g.x = __fmul_rn(g.x, a.gscale); g.y = __fmul_rn(g.y, a.gscale); g.z = __fmul_rn(g.z, a.gscale); g.w = __fmul_rn(g.w, a.gscale); }
      p.x = adam_one(g.x, m.x, v.x, p.x, lr_t, omb1, omb2, a.eps);
      p.y = adam_one(g.y, m.y, v.y, p.y, lr_t, omb1, omb2, a.eps);
      p.z = adam_one(g.z, m.z, v.z, p.z, lr_t, omb1, omb2, a.eps);
      p.w = adam_one(g.w, m.w, v.w, p.w, lr_t, omb1, omb2, a.eps);
      for (int r = 0; r < a.world; ++r) {
        if (r == a.rank) continue;
        uint4* dst = ll_pstage(a.signals[r], i);
        st_ll(dst, p.x, p.y, e);
        st_ll(dst + kLLWord1, p.z, p.w, e);
      }
      reinterpret_cast<float4*>(a.m)[i] = m; reinterpret_cast<float4*>(a.v)[i] = v;
      reinterpret_cast<float4*>(a.param)[i] = p;
    }
  }
  if (a.trace) { __syncthreads(); if (threadIdx.x == 0) atomicMax(&a.trace[2], now_ns()); }
  // ---- C. the other shards arrive in my parameter staging area
  for (int64_t i = tid; i < n4 && ok; i += nthr) {
    if (i >= lo && i < hi) continue;
    float4 p;
    ok = wait_ll(ll_pstage(me, i), e, p, abort_word);
    if (ok) reinterpret_cast<float4*>(a.param)[i] = p;
  }
  if (!ok) declare_dead(a);          // sticky: every later launch returns at entry (the epochs no longer matter)
  __syncthreads();                   // every thread of the CTA has read the counter
  if (threadIdx.x != 0) return;
  a.ll_epochs[blockIdx.x] = e;
  if (a.trace) {
    // [0] first CTA in, [1] last CTA past A, [2] past B, [3] past C; the CTA that takes the last ticket ([10]) adds the three
    // phases and the gap since the previous exchange to [4..7], steps in [8], previous end in [9]
    unsigned long long* t = a.trace;
    atomicMax(&t[3], now_ns());
    __threadfence();
    if (atomicAdd(&t[10], 1ull) != (unsigned long long)gridDim.x - 1) return;
    __threadfence();
    const unsigned long long t0 = atomicMin(&t[0], ~0ull), tA = atomicMax(&t[1], 0ull), tB = atomicMax(&t[2], 0ull), tC = atomicMax(&t[3], 0ull);
    t[4] += tA - t0; t[5] += tB - tA; t[6] += tC - tB;
    if (t[9]) t[7] += t0 - t[9];
    t[8] += 1; t[9] = tC;
    t[0] = ~0ull; t[1] = 0; t[2] = 0; t[3] = 0; t[10] = 0;
    __threadfence();
  }
}

// WF2CU_P2P_TRACE=1: the LL kernel stamps its phases with %globaltimer; wf2cu_p2p_detach prints the per-step means
static unsigned long long* g_p2p_trace = nullptr;
static unsigned long long* p2p_trace_buffer() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("WF2CU_P2P_TRACE"); on = (e && e[0] == '1') ? 1 : 0; }
  if (!on) return nullptr;
  if (!g_p2p_trace) {
    if (cudaMalloc(&g_p2p_trace, 128) != cudaSuccess) return nullptr;
    unsigned long long init[16] = {~0ull};
    cudaMemcpy(g_p2p_trace, init, 128, cudaMemcpyHostToDevice);
  }
  return g_p2p_trace;
}
static void p2p_trace_report(int rank) {
  if (!g_p2p_trace) return;
  unsigned long long t[16];
  if (cudaMemcpy(t, g_p2p_trace, 128, cudaMemcpyDeviceToHost) == cudaSuccess && t[8])
    fprintf(stderr, "[wf2cu p2p trace rank %d] %llu LL steps: push gradients %.2f us, reduce + Adam + push parameters %.2f us, receive parameters %.2f us, "
            "between exchanges %.2f us\n", rank, t[8], t[4] / 1e3 / t[8], t[5] / 1e3 / t[8], t[6] / 1e3 / t[8], t[8] > 1 ? t[7] / 1e3 / (t[8] - 1) : 0.0);
  unsigned long long init[16] = {~0ull};
  cudaMemcpy(g_p2p_trace, init, 128, cudaMemcpyHostToDevice);
}

// WF2CU_P2P_LL=0: always the pull kernel (A/B measurements)
static bool p2p_ll_enabled() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("WF2CU_P2P_LL"); on = (e && e[0] == '0') ? 0 : 1; }
  return on != 0;
}

static void p2p_close() {
  P2PState& p = g_p2p;
  for (int r = 0; r < kMaxPeers; ++r) {
    if (p.opened_grad[r]) cudaIpcCloseMemHandle(p.peer_grad_base[r]);
    if (p.opened_param[r]) cudaIpcCloseMemHandle(p.peer_param_base[r]);
    if (p.opened_signal[r]) cudaIpcCloseMemHandle(p.peer_signal[r]);
    p.opened_grad[r] = p.opened_signal[r] = p.opened_param[r] = false;
    p.peer_grad_base[r] = nullptr; p.peer_signal[r] = nullptr; p.peer_param_base[r] = nullptr;
  }
  p.rank = -1; p.world = 0; p.ll_grid = 0;
}

}  // namespace wf2

using namespace wf2;

extern "C" {

int wf2cu_p2p_export(void* dev_alloc, void* handle64) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(dev_alloc && handle64, "wf2cu_p2p_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  cudaIpcMemHandle_t h;
  WF2_CUDA(cudaIpcGetMemHandle(&h, dev_alloc));
  memcpy(handle64, &h, 64);
  return 0;
}

int wf2cu_p2p_signal_export(void* handle64) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(handle64, "wf2cu_p2p_signal_export: null argument");
  P2PState& p = g_p2p;
  if (!p.my_signal) {
    WF2_CUDA(cudaMalloc(&p.my_signal, kWindowBytes));          // signal page + the LL staging areas (epoch 0 = never written)
    WF2_CUDA(cudaMemset(p.my_signal, 0, kWindowBytes));
    WF2_CUDA(cudaMalloc(&p.local, kLocalBytes));
    WF2_CUDA(cudaMemset(p.local, 0, kLocalBytes));
    WF2_CUDA(cudaHostAlloc((void**)&p.host_flag, 64, cudaHostAllocMapped));
    *p.host_flag = 0;
    WF2_CUDA(cudaHostGetDevicePointer((void**)&p.host_flag_dev, p.host_flag, 0));
  }
  cudaIpcMemHandle_t h;
  WF2_CUDA(cudaIpcGetMemHandle(&h, p.my_signal));
  memcpy(handle64, &h, 64);
  return 0;
}

int wf2cu_p2p_attach(int rank, int world, void* local_grad_alloc, void* local_param_alloc, const void* grad_handles,
                     const void* param_handles, const void* signal_handles) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "wf2cu_p2p_attach: bad rank/world (%d/%d, at most %d peers)", rank, world, kMaxPeers);
  WF2_CHECK(local_grad_alloc && local_param_alloc && grad_handles && param_handles && signal_handles, "wf2cu_p2p_attach: null argument");
  P2PState& p = g_p2p;
  if (p.world) return set_error(WF2CU_ERR_COMM, "wf2cu_p2p_attach: already attached (call wf2cu_p2p_detach)");
  if (!p.my_signal) return set_error(WF2CU_ERR_COMM, "wf2cu_p2p_attach: call wf2cu_p2p_signal_export first");
  // The signal page and the epoch were zeroed when the page was created and by wf2cu_p2p_detach — NOT here: a faster peer
  // that has finished its own attach may already have posted ready[peer] = 1 into this page, and a memset now would erase it
  // (the first step would then end in a timeout).  Contract: a host barrier across ranks before wf2cu_p2p_detach.
  for (int r = 0; r < world; ++r) {
    if (r == rank) { p.peer_grad_base[r] = local_grad_alloc; p.peer_param_base[r] = local_param_alloc; p.peer_signal[r] = p.my_signal; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)grad_handles + 64 * r, 64);
    cudaError_t e = cudaIpcOpenMemHandle(&p.peer_grad_base[r], h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { p2p_close(); return cuda_fail(e, "cudaIpcOpenMemHandle(gradient arena)", __FILE__, __LINE__); }
    p.opened_grad[r] = true;
    memcpy(&h, (const char*)param_handles + 64 * r, 64);
    e = cudaIpcOpenMemHandle(&p.peer_param_base[r], h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { p2p_close(); return cuda_fail(e, "cudaIpcOpenMemHandle(parameter arena)", __FILE__, __LINE__); }
    p.opened_param[r] = true;
    memcpy(&h, (const char*)signal_handles + 64 * r, 64);
    void* sp = nullptr;
    e = cudaIpcOpenMemHandle(&sp, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { p2p_close(); return cuda_fail(e, "cudaIpcOpenMemHandle(signal page)", __FILE__, __LINE__); }
    p.peer_signal[r] = (SignalPage*)sp;
    p.opened_signal[r] = true;
  }
  p.rank = rank; p.world = world;
  WF2_CUDA(cudaDeviceSynchronize());
  return 0;
}

int wf2cu_p2p_allreduce_adam_step(void* param, int64_t p_off, int64_t g_off, void* m, int64_t m_off, void* v, int64_t v_off,
                                  int64_t n, float lr_t, const void* lr_t_dev, float beta1, float beta2, float eps) {
  WF2_REQUIRE_INIT();
  P2PState& p = g_p2p;
  if (!p.world) return set_error(WF2CU_ERR_COMM, "wf2cu_p2p_allreduce_adam_step: not attached (call wf2cu_p2p_attach)");
  WF2_CHECK(param && m && v && n >= 0, "wf2cu_p2p_allreduce_adam_step: bad arguments");
  WF2_CHECK(((p_off | g_off | m_off | v_off) & 3) == 0, "wf2cu_p2p_allreduce_adam_step: offsets must be multiples of 4 elements");
  if (n == 0) return 0;
  P2PArgs a;
  memset(&a, 0, sizeof(a));
  a.param = (float*)param + p_off; a.m = (float*)m + m_off; a.v = (float*)v + v_off;
  a.lr_dev = (const float*)lr_t_dev; a.lr_t = lr_t; a.b1 = beta1; a.b2 = beta2; a.eps = eps;
  a.gscale = 1.0f / (float)p.world;
  a.n = n; a.world = p.world; a.rank = p.rank;
  for (int r = 0; r < p.world; ++r) {
    a.grads[r] = (const float*)p.peer_grad_base[r] + g_off; a.params[r] = (float*)p.peer_param_base[r] + p_off;
    a.signals[r] = p.peer_signal[r];
  }
  if (a.params[p.rank] != a.param) return set_error(WF2CU_ERR_INVALID, "wf2cu_p2p_allreduce_adam_step: param is not the arena given to wf2cu_p2p_attach");
  a.local = p.local;
  a.host_flag = p.host_flag_dev;
  a.grad = a.grads[p.rank];
  if (n <= kLLMaxElems && p2p_ll_enabled()) {
    a.trace = p2p_trace_buffer();
    // LL (push) kernel: co-resident grid, every CTA pushes before it polls
    const int64_t n4ll = (n + 3) >> 2;
    int64_t g = (n4ll + 255) / 256;
    // every CTA must be resident (they wait for each other's peers): never more CTAs than the device can hold at once
    static int per_sm = 0;
    if (!per_sm) {
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, allreduce_adam_ll_kernel, 256, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
      if (per_sm > 2) per_sm = 2;
    }
    int64_t capll = (int64_t)S().sm_count * per_sm;
    if (capll > kLLMaxCtas) capll = kLLMaxCtas;
    if (g > capll) g = capll;
    if (p.ll_grid && p.ll_grid != g)
      return set_error(WF2CU_ERR_INVALID, "wf2cu_p2p_allreduce_adam_step: the bucket size changed within one attachment (detach and attach again)");
    p.ll_grid = g;
    a.ll_epochs = p.local + 64;
    WF2_CUDA(launch_pdl(allreduce_adam_ll_kernel, dim3((unsigned)g), dim3(256), 0, S().stream, a));
    return post_launch("wf2cu_p2p_allreduce_adam_step(LL)");
  }
  const int64_t n4 = (n + 3) >> 2;
  const int64_t per = (n4 + p.world - 1) / p.world;
  int64_t grid = (per + 255) / 256;
  const int64_t cap = (int64_t)S().sm_count * 4;
  if (grid > cap) grid = cap;
  allreduce_adam_kernel<<<(unsigned)grid, 256, 0, S().stream>>>(a);
  return post_launch("wf2cu_p2p_allreduce_adam_step");
}

int wf2cu_p2p_status(int* timed_out) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(timed_out, "wf2cu_p2p_status: null out");
  *timed_out = 0;
  if (!g_p2p.host_flag) return 0;
  // no sync: the kernel writes this mapped pinned word with a system-scope release store when it gives up
  *timed_out = (int)*(volatile uint32_t*)g_p2p.host_flag;
  return 0;
}

int wf2cu_p2p_detach(void) {
  if (g_p2p.world) {
    cudaStreamSynchronize(S().stream);
    p2p_trace_report(g_p2p.rank);
    p2p_close();
    // every rank's kernels have finished (the caller's barrier): a fresh attachment starts at epoch 0 with clean flags
    if (g_p2p.my_signal) cudaMemset(g_p2p.my_signal, 0, kWindowBytes);
    if (g_p2p.local) cudaMemset(g_p2p.local, 0, kLocalBytes);
    if (g_p2p.host_flag) *g_p2p.host_flag = 0;
  }
  return 0;
}

}  // extern "C"
