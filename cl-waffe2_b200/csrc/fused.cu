// fused.cu — fused element-wise chains (`wf2cu_fused`): one pass over equal-shape views running a
// small register program.  This is what the backend's `on-finalizing-compiling` pass emits for the
// composites that have no node of their own in the reference: `!sigmoid`
// (`source/nn/activation.lisp:43`), `!relu` and its MulNode backward (`:11-26`), `!gelu` (`:61-66`), the MSE
// difference chain (`nn/criterion.lisp:73-79`), the element-wise part of softmax-cross-entropy and
// `sce-diff-static` (`:81-117`), SGD on views (`optimizers/impls/sgd.lisp:20-26`), and chains of unary math
// nodes such as the reference's own `sin(sin(sin x))` benchmark (`benchmarks/compose.lisp:21-26`).
//
// Two executors with the same per-element operation order and roundings (no FMA contraction anywhere):
//   * COMPILED: the programs the host fusion passes emit are listed below as compile-time tables; for each
//     one `fused_cat_kernel<P, V>` is the program unrolled at compile time — every register index is a
//     constant, so the program's registers are machine registers, each thread keeps V 128-bit vectors per
//     view in flight, and the kernel is HBM-bound like `ew_vec_kernel`.  A program handed to wf2cu_fused is
//     matched field by field (immediates are run-time arguments).
//   * INTERPRETED: anything else (unknown programs, strided / misaligned views) runs on the per-element
//     opcode interpreter.  Its register file is indexed at run time, lives in local memory and reaches only
//     ~0.14 of the HBM peak (measured round 1): it is the fallback, not the product path.
#include "common.cuh"
#include "math.cuh"

namespace wf2 {

constexpr int kMaxInstr = 64;
constexpr int kMaxRegs = 16;

static int g_fused_mode = 0;        // 0 auto, 1 force the interpreter (parity tests compare the two executors)
static int g_fused_last_path = 0;   // 0 none yet, 1 interpreter, 2 compiled catalogue kernel

// ------------------------------------------------------------------ interpreter (fallback)
struct FusedArgs {
  int n_instr;
  int nviews;
  int4 prog[kMaxInstr];
  float* base[kMaxOperands];
  int rank;
  int64_t shape[kMaxRank];
  int64_t stride[kMaxOperands][kMaxRank];
  int64_t numel;
  int vec;  // all views contiguous (stride 1 or 0 over a rank-1 plan) and 16-byte aligned
};

__device__ __forceinline__ float fused_eval(const int4 ins, const float* r) {
  const int opc = ins.x & 0xff;
  switch (opc) {
    case WF2CU_P_IMM: return __int_as_float(ins.w);
    case WF2CU_P_ADD: return __fadd_rn(r[ins.z], r[ins.w]);
    case WF2CU_P_SUB: return __fsub_rn(r[ins.z], r[ins.w]);
    case WF2CU_P_MUL: return __fmul_rn(r[ins.z], r[ins.w]);
    case WF2CU_P_DIV: return __fdiv_rn(r[ins.z], r[ins.w]);
    case WF2CU_P_UNARY: return unary_dispatch(ins.w, r[ins.z]);
    case WF2CU_P_POW: return pow_ref(r[ins.z], r[ins.w]);
    case WF2CU_P_CMPSEL: return cmp_dispatch(ins.x >> 8, r[ins.z], r[ins.w]) ? 1.0f : 0.0f;
    case WF2CU_P_MOV: return r[ins.z];
    default: return 0.f;
  }
}

__global__ void __launch_bounds__(256) fused_kernel(const __grid_constant__ FusedArgs a) {
  pdl_prologue();
  const int64_t gstride = (int64_t)gridDim.x * 256;
  if (a.vec) {
    const int64_t n4 = a.numel >> 2;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += gstride) {
      float r[4][kMaxRegs];
      for (int k = 0; k < a.n_instr; ++k) {
        const int4 ins = a.prog[k];
        const int opc = ins.x & 0xff;
        if (opc == WF2CU_P_LOAD) {
          if (a.stride[ins.z][0] == 0) {
            const float s = a.base[ins.z][0];
            r[0][ins.y] = s; r[1][ins.y] = s; r[2][ins.y] = s; r[3][ins.y] = s;
          } else {
            const float4 v = reinterpret_cast<const float4*>(a.base[ins.z])[i];
            r[0][ins.y] = v.x; r[1][ins.y] = v.y; r[2][ins.y] = v.z; r[3][ins.y] = v.w;
          }
        } else if (opc == WF2CU_P_STORE) {
          reinterpret_cast<float4*>(a.base[ins.z])[i] = make_float4(r[0][ins.y], r[1][ins.y], r[2][ins.y], r[3][ins.y]);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) r[e][ins.y] = fused_eval(ins, r[e]);
        }
      }
    }
    return;
  }
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < a.numel; idx += gstride) {
    int64_t off[kMaxOperands];
    for (int j = 0; j < a.nviews; ++j) off[j] = 0;
    int64_t rem = idx;
    for (int d = a.rank - 1; d >= 0; --d) {
      const int64_t q = rem / a.shape[d], rr = rem - q * a.shape[d];
      rem = q;
      for (int j = 0; j < a.nviews; ++j) off[j] += rr * a.stride[j][d];
    }
    float r[kMaxRegs];
    for (int k = 0; k < a.n_instr; ++k) {
      const int4 ins = a.prog[k];
      const int opc = ins.x & 0xff;
      if (opc == WF2CU_P_LOAD) r[ins.y] = a.base[ins.z][off[ins.z]];
      else if (opc == WF2CU_P_STORE) a.base[ins.z][off[ins.z]] = r[ins.y];
      else r[ins.y] = fused_eval(ins, r);
    }
  }
}

// ------------------------------------------------------------------ compiled catalogue
// A program is a type: n instructions x {opcode, dst, a, b}; `at(k, f)` is a constant expression, so after the
// recursion below is instantiated no table exists in the kernel — only its straight-line code.  IMM operands are
// wildcards here (field 3 = 0) and arrive in CatArgs::imm[k]; CMPSEL carries its compare in bits 8.. of the opcode.
constexpr int kCatMaxInstr = 16;
constexpr int kCatRegs = 4;
constexpr int kCatViews = 4;

#define L WF2CU_P_LOAD
#define ST WF2CU_P_STORE
#define IM WF2CU_P_IMM
#define AD WF2CU_P_ADD
#define SU WF2CU_P_SUB
#define MU WF2CU_P_MUL
#define DI WF2CU_P_DIV
#define UN WF2CU_P_UNARY
#define GT (WF2CU_P_CMPSEL | (WF2CU_GT << 8))
#define WF2_PROG(NAME, NVIEWS, N, ...)                                                    \
  struct NAME {                                                                           \
    static constexpr int n = N, nviews = NVIEWS;                                          \
    static __host__ __device__ constexpr int at(int k, int f) {                           \
      constexpr int t[N][4] = __VA_ARGS__;                                                \
      return t[k][f];                                                                     \
    }                                                                                     \
    static const char* name() { return #NAME; }                                           \
  };

// ops.py `sigmoid`: 1 / (1 + exp(-x))                      (nn/activation.lisp:43)
WF2_PROG(PSigmoid, 2, 7, {{L, 0, 0, 0}, {UN, 1, 0, WF2CU_NEG}, {UN, 1, 1, WF2CU_EXP}, {IM, 2, 0, 0}, {AD, 1, 2, 1}, {DI, 1, 2, 1}, {ST, 1, 1, 0}})
// ops.py `relu`: x * (x > 0 ? 1 : 0)                       (nn/activation.lisp:11-26)
WF2_PROG(PRelu, 2, 5, {{L, 0, 0, 0}, {IM, 1, 0, 0}, {GT, 2, 0, 1}, {MU, 2, 0, 2}, {ST, 2, 1, 0}})
// ops.py `gelu`: 0.5 x (1 + tanh(c2 (x + c1 x^3)))         (nn/activation.lisp:61-66, examples/gpt-2/utils.lisp:129-137)
WF2_PROG(PGelu, 2, 15, {{L, 0, 0, 0}, {MU, 1, 0, 0}, {MU, 1, 1, 0}, {IM, 2, 0, 0}, {MU, 1, 1, 2}, {AD, 1, 0, 1}, {IM, 2, 0, 0}, {MU, 1, 2, 1},
                        {UN, 1, 1, WF2CU_TANH}, {IM, 2, 0, 0}, {AD, 1, 2, 1}, {IM, 2, 0, 0}, {MU, 2, 2, 0}, {MU, 1, 2, 1}, {ST, 1, 1, 0}})
// nn.py `_mask_mul`: g <- g * (out > 0)                    (the `!relu` backward applied to a finished gradient)
WF2_PROG(PMaskMul, 2, 6, {{L, 0, 0, 0}, {IM, 1, 0, 0}, {GT, 1, 0, 1}, {L, 2, 1, 0}, {MU, 1, 2, 1}, {ST, 1, 1, 0}})
// nn.py `relu` backward: g <- [g +] dout * (x > 0)         views x, dout, g
WF2_PROG(PReluBw, 3, 6, {{L, 0, 0, 0}, {IM, 1, 0, 0}, {GT, 1, 0, 1}, {L, 2, 1, 0}, {MU, 1, 2, 1}, {ST, 1, 2, 0}})
WF2_PROG(PReluBwAcc, 3, 8, {{L, 0, 0, 0}, {IM, 1, 0, 0}, {GT, 1, 0, 1}, {L, 2, 1, 0}, {MU, 1, 2, 1}, {L, 3, 2, 0}, {AD, 1, 3, 1}, {ST, 1, 2, 0}})
// nn.py `sigmoid` backward: g <- [g +] dout * (s * (1 - s))    views s, dout, g
WF2_PROG(PSigmoidBw, 3, 7, {{L, 0, 0, 0}, {IM, 1, 0, 0}, {SU, 1, 1, 0}, {MU, 1, 0, 1}, {L, 2, 1, 0}, {MU, 1, 2, 1}, {ST, 1, 2, 0}})
WF2_PROG(PSigmoidBwAcc, 3, 9, {{L, 0, 0, 0}, {IM, 1, 0, 0}, {SU, 1, 1, 0}, {MU, 1, 0, 1}, {L, 2, 1, 0}, {MU, 1, 2, 1}, {L, 3, 2, 0}, {AD, 1, 3, 1}, {ST, 1, 2, 0}})
// nn.py `softmax_cross_entropy`: out = -1 * (labels * log(z + delta))   (nn/criterion.lisp:89-117)   views z, labels, out
WF2_PROG(PCrossEntropy, 3, 9, {{L, 0, 0, 0}, {IM, 1, 0, 0}, {AD, 0, 0, 1}, {UN, 0, 0, WF2CU_LOGE}, {L, 2, 1, 0}, {MU, 0, 2, 0}, {IM, 1, 0, 0}, {MU, 0, 1, 0}, {ST, 0, 2, 0}})
// nn.py `sce-diff-static`: g <- [g +] (dy * (z - labels)) / batch   (nn/criterion.lisp:81-87)   views z, labels, dy, g
WF2_PROG(PSceDiff, 4, 8, {{L, 0, 0, 0}, {L, 1, 1, 0}, {SU, 0, 0, 1}, {L, 2, 2, 0}, {MU, 0, 2, 0}, {IM, 1, 0, 0}, {DI, 0, 0, 1}, {ST, 0, 3, 0}})
WF2_PROG(PSceDiffAcc, 4, 10, {{L, 0, 0, 0}, {L, 1, 1, 0}, {SU, 0, 0, 1}, {L, 2, 2, 0}, {MU, 0, 2, 0}, {IM, 1, 0, 0}, {DI, 0, 0, 1}, {L, 3, 3, 0}, {AD, 0, 3, 0}, {ST, 0, 3, 0}})
// nn.py `mse` backward: gx <- [gx +] (d*l + d*l) ; gy <- [gy +] -(d*l + d*l)   (nn/criterion.lisp:73-79)   views l, d, g
WF2_PROG(PMseBwX, 3, 5, {{L, 0, 0, 0}, {L, 1, 1, 0}, {MU, 0, 1, 0}, {AD, 0, 0, 0}, {ST, 0, 2, 0}})
WF2_PROG(PMseBwXAcc, 3, 7, {{L, 0, 0, 0}, {L, 1, 1, 0}, {MU, 0, 1, 0}, {AD, 0, 0, 0}, {L, 2, 2, 0}, {AD, 0, 2, 0}, {ST, 0, 2, 0}})
WF2_PROG(PMseBwY, 3, 6, {{L, 0, 0, 0}, {L, 1, 1, 0}, {MU, 0, 1, 0}, {AD, 0, 0, 0}, {UN, 0, 0, WF2CU_NEG}, {ST, 0, 2, 0}})
WF2_PROG(PMseBwYAcc, 3, 8, {{L, 0, 0, 0}, {L, 1, 1, 0}, {MU, 0, 1, 0}, {AD, 0, 0, 0}, {UN, 0, 0, WF2CU_NEG}, {L, 2, 2, 0}, {AD, 0, 2, 0}, {ST, 0, 2, 0}})
// `mse` forward difference chain: out = (x - y) * (x - y)   (nn/criterion.lisp:73-79)   views x, y, out
WF2_PROG(PMseDiff, 3, 5, {{L, 0, 0, 0}, {L, 1, 1, 0}, {SU, 0, 0, 1}, {MU, 0, 0, 0}, {ST, 0, 2, 0}})
// the reference's composed benchmark: sin(sin(sin x))       (benchmarks/compose.lisp:21-26)
WF2_PROG(PSin3, 2, 5, {{L, 0, 0, 0}, {UN, 0, 0, WF2CU_SIN}, {UN, 0, 0, WF2CU_SIN}, {UN, 0, 0, WF2CU_SIN}, {ST, 0, 1, 0}})
// SGD on views: p <- p - (g * lr)                           (optimizers/impls/sgd.lisp:20-26)   views p, g
WF2_PROG(PSgd, 2, 6, {{L, 0, 0, 0}, {L, 1, 1, 0}, {IM, 2, 0, 0}, {MU, 1, 1, 2}, {SU, 0, 0, 1}, {ST, 0, 0, 0}})

#undef L
#undef ST
#undef IM
#undef AD
#undef SU
#undef MU
#undef DI
#undef UN
#undef GT

struct CatArgs {
  float* base[kCatViews];
  uint32_t bcast;             // bit j: view j is one broadcast scalar (stride 0)
  uint32_t n4;                // float4 vectors (< 2^31, checked by the dispatcher)
  float imm[kCatMaxInstr];    // imm[k]: the operand of instruction k when it is an IMM
};

// one instruction of program P applied to the thread's E = 4 V elements; everything indexed by constants
template <class P, int K, int V>
struct CatStep {
  static __device__ __forceinline__ void run(float (&r)[kCatRegs][4 * V], const CatArgs& a, const uint32_t (&idx)[V], const bool (&ok)[V]) {
    if constexpr (K < P::n) {
      constexpr int op = P::at(K, 0), opc = op & 0xff, d = P::at(K, 1), x = P::at(K, 2), y = P::at(K, 3);
      if constexpr (opc == WF2CU_P_LOAD) {
        if (a.bcast & (1u << x)) {
          const float s = a.base[x][0];
#pragma unroll
          for (int e = 0; e < 4 * V; ++e) r[d][e] = s;
        } else {
#pragma unroll
          for (int v = 0; v < V; ++v)
            if (ok[v]) {
              const float4 q = reinterpret_cast<const float4*>(a.base[x])[idx[v]];
              r[d][4 * v] = q.x; r[d][4 * v + 1] = q.y; r[d][4 * v + 2] = q.z; r[d][4 * v + 3] = q.w;
            }
        }
      } else if constexpr (opc == WF2CU_P_STORE) {
#pragma unroll
        for (int v = 0; v < V; ++v)
          if (ok[v])
            reinterpret_cast<float4*>(a.base[x])[idx[v]] = make_float4(r[d][4 * v], r[d][4 * v + 1], r[d][4 * v + 2], r[d][4 * v + 3]);
      } else {
#pragma unroll
        for (int e = 0; e < 4 * V; ++e) {
          if constexpr (opc == WF2CU_P_IMM) r[d][e] = a.imm[K];
          else if constexpr (opc == WF2CU_P_ADD) r[d][e] = __fadd_rn(r[x][e], r[y][e]);
          else if constexpr (opc == WF2CU_P_SUB) r[d][e] = __fsub_rn(r[x][e], r[y][e]);
          else if constexpr (opc == WF2CU_P_MUL) r[d][e] = __fmul_rn(r[x][e], r[y][e]);
          else if constexpr (opc == WF2CU_P_DIV) r[d][e] = __fdiv_rn(r[x][e], r[y][e]);
          else if constexpr (opc == WF2CU_P_UNARY) r[d][e] = Unary<y>::apply(r[x][e]);
          else if constexpr (opc == WF2CU_P_POW) r[d][e] = pow_ref(r[x][e], r[y][e]);
          else if constexpr (opc == WF2CU_P_CMPSEL) r[d][e] = cmp_dispatch(op >> 8, r[x][e], r[y][e]) ? 1.0f : 0.0f;
          else if constexpr (opc == WF2CU_P_MOV) r[d][e] = r[x][e];
        }
      }
      CatStep<P, K + 1, V>::run(r, a, idx, ok);
    }
  }
};

// One CTA = 256 threads x V vectors = 1024 V consecutive floats; thread t owns vectors t, t + 256, ... (coalesced).
// Views may alias (in-place programs): every element is read before it is written, by the same thread.
template <class P, int V>
__global__ void __launch_bounds__(256) fused_cat_kernel(const __grid_constant__ CatArgs a) {
  pdl_prologue();
  const uint32_t first = blockIdx.x * (uint32_t)(256 * V) + threadIdx.x;
  uint32_t idx[V];
  bool ok[V];
#pragma unroll
  for (int v = 0; v < V; ++v) { idx[v] = first + (uint32_t)v * 256; ok[v] = idx[v] < a.n4; }
  float r[kCatRegs][4 * V];
#pragma unroll
  for (int q = 0; q < kCatRegs; ++q)
#pragma unroll
    for (int e = 0; e < 4 * V; ++e) r[q][e] = 0.f;
  CatStep<P, 0, V>::run(r, a, idx, ok);
}

template <class P>
static bool cat_match(const int32_t* program, int n_instr, int nviews, CatArgs* a) {
  if (n_instr != P::n || nviews != P::nviews) return false;
  for (int k = 0; k < P::n; ++k) {
    const int32_t* q = program + 4 * k;
    if (q[0] != P::at(k, 0) || q[1] != P::at(k, 1)) return false;
    const int opc = q[0] & 0xff;
    if (opc == WF2CU_P_IMM) { memcpy(&a->imm[k], &q[3], 4); continue; }
    if (q[2] != P::at(k, 2)) return false;
    if (opc == WF2CU_P_LOAD || opc == WF2CU_P_STORE || opc == WF2CU_P_MOV) continue;
    if (q[3] != P::at(k, 3)) return false;
  }
  return true;
}

template <class P>
static int cat_launch(const CatArgs& a) {
  static_assert(P::n <= kCatMaxInstr && P::nviews <= kCatViews, "catalogue program too large");
  constexpr int V = P::nviews >= 4 ? 2 : 4;      // <= 12 128-bit loads in flight per thread
  const uint32_t per = 256 * V;
  const uint32_t grid = (a.n4 + per - 1) / per;
  WF2_CUDA(launch_pdl(fused_cat_kernel<P, V>, dim3(grid), dim3(256), 0, S().stream, a));
  return post_launch(P::name());
}

template <class... Ps> struct Catalogue;
template <> struct Catalogue<> {
  static int try_launch(const int32_t*, int, int, CatArgs*, bool*) { return 0; }
  static int count() { return 0; }
};
template <class P, class... Rest> struct Catalogue<P, Rest...> {
  static int try_launch(const int32_t* program, int n_instr, int nviews, CatArgs* a, bool* taken) {
    if (cat_match<P>(program, n_instr, nviews, a)) { *taken = true; return cat_launch<P>(*a); }
    return Catalogue<Rest...>::try_launch(program, n_instr, nviews, a, taken);
  }
  static int count() { return 1 + Catalogue<Rest...>::count(); }
};
using TheCatalogue = Catalogue<PSigmoid, PRelu, PGelu, PMaskMul, PReluBw, PReluBwAcc, PSigmoidBw, PSigmoidBwAcc, PCrossEntropy,
                               PSceDiff, PSceDiffAcc, PMseBwX, PMseBwXAcc, PMseBwY, PMseBwYAcc, PMseDiff, PSin3, PSgd>;

}  // namespace wf2

using namespace wf2;

extern "C" {

int wf2cu_fused_set_mode(int mode) {
  WF2_CHECK(mode == 0 || mode == 1, "wf2cu_fused_set_mode: mode %d (0 auto, 1 force the interpreter)", mode);
  g_fused_mode = mode;
  return 0;
}

int wf2cu_fused_last_path(int* out) {
  WF2_CHECK(out, "wf2cu_fused_last_path: null out");
  *out = g_fused_last_path;
  return 0;
}

int wf2cu_fused_catalogue_size(int* out) {
  WF2_CHECK(out, "wf2cu_fused_catalogue_size: null out");
  *out = TheCatalogue::count();
  return 0;
}

int wf2cu_fused(const int32_t* program, int n_instr, const wf2cu_view* views, int nviews) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(program && views, "wf2cu_fused: null argument");
  WF2_CHECK(n_instr >= 1 && n_instr <= kMaxInstr, "wf2cu_fused: %d instructions (max %d)", n_instr, kMaxInstr);
  WF2_CHECK(nviews >= 1 && nviews <= kMaxOperands, "wf2cu_fused: %d views (max %d)", nviews, kMaxOperands);
  const wf2cu_view* vp[kMaxOperands];
  for (int j = 0; j < nviews; ++j) vp[j] = &views[j];
  Plan p;
  int rc = build_plan(vp, nviews, &p, "wf2cu_fused");
  if (rc) return rc;
  if (p.numel == 0) return 0;
  FusedArgs a;
  memset(&a, 0, sizeof(a));
  a.n_instr = n_instr; a.nviews = nviews;
  bool stored[kMaxOperands] = {false};
  for (int k = 0; k < n_instr; ++k) {
    const int32_t* q = program + 4 * k;
    const int opc = q[0] & 0xff;
    WF2_CHECK(opc >= WF2CU_P_LOAD && opc <= WF2CU_P_MOV, "wf2cu_fused: bad opcode %d at %d", opc, k);
    WF2_CHECK(q[1] >= 0 && q[1] < kMaxRegs, "wf2cu_fused: bad dst register at %d", k);
    if (opc == WF2CU_P_LOAD || opc == WF2CU_P_STORE) {
      WF2_CHECK(q[2] >= 0 && q[2] < nviews, "wf2cu_fused: bad view index at %d", k);
      if (opc == WF2CU_P_STORE) stored[q[2]] = true;
    } else if (opc != WF2CU_P_IMM) {
      WF2_CHECK(q[2] >= 0 && q[2] < kMaxRegs, "wf2cu_fused: bad source register at %d", k);
      if (opc != WF2CU_P_UNARY && opc != WF2CU_P_MOV)
        WF2_CHECK(q[3] >= 0 && q[3] < kMaxRegs, "wf2cu_fused: bad source register at %d", k);
      if (opc == WF2CU_P_UNARY)
        WF2_CHECK(q[3] >= 0 && q[3] < WF2CU_UNARY_COUNT, "wf2cu_fused: bad unary function %d at %d", q[3], k);
    }
    a.prog[k] = make_int4(q[0], q[1], q[2], q[3]);
  }
  bool vec = (p.rank == 1) && (p.numel % 4 == 0);
  for (int j = 0; j < nviews; ++j) {
    a.base[j] = fptr(&views[j]);
    for (int d = 0; d < p.rank; ++d) {
      a.stride[j][d] = p.stride[j][d];
      if (stored[j] && p.stride[j][d] == 0 && p.shape[d] > 1)
        return set_error(WF2CU_ERR_INVALID, "wf2cu_fused: stored view %d has a broadcast axis", j);
    }
    const int64_t s0 = p.stride[j][0];
    if (!(s0 == 1 || s0 == 0)) vec = false;
    if (s0 == 1 && (reinterpret_cast<uintptr_t>(a.base[j]) & 15)) vec = false;
  }
  a.rank = p.rank;
  for (int d = 0; d < p.rank; ++d) a.shape[d] = p.shape[d];
  a.numel = p.numel;
  a.vec = vec ? 1 : 0;
  // ---- compiled catalogue kernel: dense, aligned views and a program the host passes are known to emit
  if (vec && g_fused_mode == 0 && nviews <= kCatViews && n_instr <= kCatMaxInstr && p.numel / 4 < (int64_t(1) << 31)) {
    CatArgs c;
    memset(&c, 0, sizeof(c));
    for (int j = 0; j < nviews; ++j) {
      c.base[j] = a.base[j];
      if (p.stride[j][0] == 0) c.bcast |= (1u << j);
    }
    c.n4 = (uint32_t)(p.numel / 4);
    bool taken = false;
    rc = TheCatalogue::try_launch(program, n_instr, nviews, &c, &taken);
    if (taken) { g_fused_last_path = 2; return rc; }
  }
  const int64_t work = vec ? p.numel / 4 : p.numel;
  int64_t grid = (work + 255) / 256;
  const int64_t cap = (int64_t)S().sm_count * 8;
  if (grid > cap) grid = cap;
  WF2_CUDA(launch_pdl(fused_kernel, dim3((unsigned)grid), dim3(256), 0, S().stream, a));
  g_fused_last_path = 1;
  return post_launch("wf2cu_fused");
}

}  // extern "C"
