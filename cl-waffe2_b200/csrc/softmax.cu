// softmax.cu — `!softmax` as ONE pass over HBM.
//
// The reference has no softmax node; `(!softmax x :avoid-overflow t :axis 1)` is the composite
//     x1 = x - mean(x, axis) ; e = exp(x1) ; z = sum(e, axis) ; e / z
// (`source/nn/activation.lisp:126-159`; note: MEAN-shift, not max-shift), or with
// `:avoid-overflow nil`:  e = exp(x) ; e / sum(e, axis)  (used by softmax-cross-entropy,
// `source/nn/criterion.lisp:163`).  That is >= 8 full memory passes on the CPU path (two disguised
// reductions + 5 element-wise nodes).  Here a row lives in registers between the three phases,
// so traffic is the algorithmic minimum: one read + one write.
//
// The arithmetic keeps the reference's operation order per element: mean = sum/n (true divide),
// x - mean, expf, e / z (true divide); only the two sums are trees instead of sequential.
#include "common.cuh"
#include "math.cuh"

namespace wf2 {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

template <int THREADS>
__device__ __forceinline__ float block_sum(float v, float* smem) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) smem[w] = v;
  __syncthreads();
  float t = (lane < THREADS / 32) ? smem[lane] : 0.f;
  t = warp_sum(t);   // every warp computes the same total
  __syncthreads();
  return t;
}

struct SmRowArgs {
  const float* x;
  float* out;
  int nouter;                 // 0..4 outer axes
  uint32_t oshape[4];
  int64_t ostride_x[4], ostride_o[4];
  int64_t cols;
  uint32_t rows;
  int avoid_overflow;
};

__device__ __forceinline__ void decode_row(const SmRowArgs& a, uint32_t row, int64_t& xo, int64_t& oo) {
  xo = 0; oo = 0;
  for (int d = a.nouter - 1; d >= 0; --d) {
    const uint32_t q = row / a.oshape[d];
    const uint32_t r = row - q * a.oshape[d];
    row = q;
    xo += (int64_t)r * a.ostride_x[d];
    oo += (int64_t)r * a.ostride_o[d];
  }
}

// one CTA per row, row cached in registers as float4s: cols <= THREADS * 4 * VPT
template <int THREADS, int VPT>
__global__ void __launch_bounds__(THREADS) softmax_row_block_kernel(const __grid_constant__ SmRowArgs a) {
  __shared__ float smem[THREADS / 32];
  pdl_prologue();
  int64_t xo, oo;
  decode_row(a, blockIdx.x, xo, oo);
  const float4* xp = reinterpret_cast<const float4*>(a.x + xo);
  float4* op = reinterpret_cast<float4*>(a.out + oo);
  const int64_t n4 = a.cols >> 2;
  float4 v[VPT];
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int64_t i = threadIdx.x + (int64_t)k * THREADS;
    v[k] = (i < n4) ? xp[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  asm volatile("" ::: "memory");   // all row loads in flight before the first use
  if (a.avoid_overflow) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; ++k) s = __fadd_rn(s, __fadd_rn(__fadd_rn(v[k].x, v[k].y), __fadd_rn(v[k].z, v[k].w)));
    s = block_sum<THREADS>(s, smem);
    const float mean = __fdiv_rn(s, (float)a.cols);
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      v[k].x = __fsub_rn(v[k].x, mean); v[k].y = __fsub_rn(v[k].y, mean);
      v[k].z = __fsub_rn(v[k].z, mean); v[k].w = __fsub_rn(v[k].w, mean);
    }
  }
  float z = 0.f;
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int64_t i = threadIdx.x + (int64_t)k * THREADS;
    if (i < n4) {
      v[k].x = expf(v[k].x); v[k].y = expf(v[k].y); v[k].z = expf(v[k].z); v[k].w = expf(v[k].w);
      z = __fadd_rn(z, __fadd_rn(__fadd_rn(v[k].x, v[k].y), __fadd_rn(v[k].z, v[k].w)));
    }
  }
  z = block_sum<THREADS>(z, smem);
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int64_t i = threadIdx.x + (int64_t)k * THREADS;
    if (i < n4) {
      float4 r;
      r.x = __fdiv_rn(v[k].x, z); r.y = __fdiv_rn(v[k].y, z); r.z = __fdiv_rn(v[k].z, z); r.w = __fdiv_rn(v[k].w, z);
      op[i] = r;
    }
  }
}

// one warp per row, scalar register cache: cols <= 1024, any alignment, strided inner axis allowed
struct SmWarpExtra { int64_t sx, so; };
__global__ void __launch_bounds__(256) softmax_row_warp_kernel(const __grid_constant__ SmRowArgs a, const SmWarpExtra e) {
  const uint32_t row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= a.rows) return;
  const int lane = threadIdx.x & 31;
  int64_t xo, oo;
  decode_row(a, row, xo, oo);
  const float* xp = a.x + xo;
  float* op = a.out + oo;
  float v[32];
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    const int64_t c = lane + 32 * k;
    v[k] = (c < a.cols) ? xp[c * e.sx] : 0.f;
  }
  if (a.avoid_overflow) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) s = __fadd_rn(s, v[k]);
    s = warp_sum(s);
    const float mean = __fdiv_rn(s, (float)a.cols);
#pragma unroll
    for (int k = 0; k < 32; ++k) v[k] = __fsub_rn(v[k], mean);
  }
  float z = 0.f;
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    const int64_t c = lane + 32 * k;
    if (c < a.cols) { v[k] = expf(v[k]); z = __fadd_rn(z, v[k]); }
  }
  z = warp_sum(z);
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    const int64_t c = lane + 32 * k;
    if (c < a.cols) op[c * e.so] = __fdiv_rn(v[k], z);
  }
}

// long rows: three passes over global/L2 (row does not fit the register cache)
__global__ void __launch_bounds__(512) softmax_row_stream_kernel(const __grid_constant__ SmRowArgs a, const SmWarpExtra e) {
  __shared__ float smem[16];
  int64_t xo, oo;
  decode_row(a, blockIdx.x, xo, oo);
  const float* xp = a.x + xo;
  float* op = a.out + oo;
  float mean = 0.f;
  if (a.avoid_overflow) {
    float s = 0.f;
    for (int64_t c = threadIdx.x; c < a.cols; c += 512) s = __fadd_rn(s, xp[c * e.sx]);
    s = block_sum<512>(s, smem);
    mean = __fdiv_rn(s, (float)a.cols);
  }
  float z = 0.f;
  for (int64_t c = threadIdx.x; c < a.cols; c += 512) {
    float t = xp[c * e.sx];
    if (a.avoid_overflow) t = __fsub_rn(t, mean);
    z = __fadd_rn(z, expf(t));
  }
  z = block_sum<512>(z, smem);
  for (int64_t c = threadIdx.x; c < a.cols; c += 512) {
    float t = xp[c * e.sx];
    if (a.avoid_overflow) t = __fsub_rn(t, mean);
    op[c * e.so] = __fdiv_rn(expf(t), z);
  }
}

// softmax over a NON-innermost axis: one thread per (outer, inner) lane, loops over R.
// Adjacent threads take adjacent inner elements, so loads coalesce when the inner stride is 1.
struct SmLaneArgs {
  const float* x;
  float* out;
  int nlane;                       // lane axes (all axes except the softmax axis), <= 7
  int64_t lshape[kMaxRank], lstride_x[kMaxRank], lstride_o[kMaxRank];
  int64_t R, sR_x, sR_o;
  int64_t nlanes;
  int avoid_overflow;
  float pre_div;                   // != 0: every element is divided by it first (the `(!div w n)` node in front, folded in)
  float pre_mul;                   // != 0: pre_div is a power of two; x * pre_mul is then the same real number as
                                   // x / pre_div, so the correctly rounded product IS the quotient (1 instruction, not ~10)
};

template <int RCACHE>
__global__ void __launch_bounds__(256) softmax_lane_kernel(const __grid_constant__ SmLaneArgs a) {
  const int64_t gstride = (int64_t)gridDim.x * 256;
  for (int64_t l = (int64_t)blockIdx.x * 256 + threadIdx.x; l < a.nlanes; l += gstride) {
    int64_t rem = l, xo = 0, oo = 0;
    for (int d = a.nlane - 1; d >= 0; --d) {
      const int64_t q = rem / a.lshape[d], r = rem - q * a.lshape[d];
      rem = q; xo += r * a.lstride_x[d]; oo += r * a.lstride_o[d];
    }
    const float* xp = a.x + xo;
    float* op = a.out + oo;
    if (RCACHE > 0) {
      float v[RCACHE > 0 ? RCACHE : 1];
#pragma unroll
      for (int k = 0; k < RCACHE; ++k) v[k] = (k < a.R) ? xp[k * a.sR_x] : 0.f;
      if (a.pre_div != 0.f) {
#pragma unroll
        for (int k = 0; k < RCACHE; ++k) v[k] = __fdiv_rn(v[k], a.pre_div);
      }
      if (a.avoid_overflow) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < RCACHE; ++k) if (k < a.R) s = __fadd_rn(s, v[k]);
        const float mean = __fdiv_rn(s, (float)a.R);
#pragma unroll
        for (int k = 0; k < RCACHE; ++k) v[k] = __fsub_rn(v[k], mean);
      }
      float z = 0.f;
#pragma unroll
      for (int k = 0; k < RCACHE; ++k) if (k < a.R) { v[k] = expf(v[k]); z = __fadd_rn(z, v[k]); }
#pragma unroll
      for (int k = 0; k < RCACHE; ++k) if (k < a.R) op[k * a.sR_o] = __fdiv_rn(v[k], z);
    } else {
      float mean = 0.f;
      if (a.avoid_overflow) {
        float s = 0.f;
        for (int64_t k = 0; k < a.R; ++k) { float t = xp[k * a.sR_x]; if (a.pre_div != 0.f) t = __fdiv_rn(t, a.pre_div); s = __fadd_rn(s, t); }
        mean = __fdiv_rn(s, (float)a.R);
      }
      float z = 0.f;
      for (int64_t k = 0; k < a.R; ++k) {
        float t = xp[k * a.sR_x];
        if (a.pre_div != 0.f) t = __fdiv_rn(t, a.pre_div);
        if (a.avoid_overflow) t = __fsub_rn(t, mean);
        z = __fadd_rn(z, expf(t));
      }
      for (int64_t k = 0; k < a.R; ++k) {
        float t = xp[k * a.sR_x];
        if (a.pre_div != 0.f) t = __fdiv_rn(t, a.pre_div);
        if (a.avoid_overflow) t = __fsub_rn(t, mean);
        op[k * a.sR_o] = __fdiv_rn(expf(t), z);
      }
    }
  }
}

// lane kernel, 4 adjacent lanes per thread: the innermost lane axis is contiguous in x and out, so a thread
// owns a float4 of lanes and walks the softmax axis with 128-bit accesses (GPT-2's softmax over the HEADS axis
// of a (8 12 1024 1024) tensor, examples/gpt-2/utils.lisp:236-242: 805 MB per call, 389 us with the scalar
// kernel above = 2.1 TB/s).  The host passes the lane decode with the innermost extent / 4 and stride 4.
template <int RCACHE>
__global__ void __launch_bounds__(256) softmax_lane_vec_kernel(const __grid_constant__ SmLaneArgs a) {
  pdl_prologue();
  const int64_t gstride = (int64_t)gridDim.x * 256;
  for (int64_t l = (int64_t)blockIdx.x * 256 + threadIdx.x; l < a.nlanes; l += gstride) {
    int64_t rem = l, xo = 0, oo = 0;
    for (int d = a.nlane - 1; d >= 0; --d) {
      const int64_t q = rem / a.lshape[d], r = rem - q * a.lshape[d];
      rem = q; xo += r * a.lstride_x[d]; oo += r * a.lstride_o[d];
    }
    const float* xp = a.x + xo;
    float* op = a.out + oo;
    float4 v[RCACHE];
#pragma unroll
    for (int k = 0; k < RCACHE; ++k)
      v[k] = (k < a.R) ? *reinterpret_cast<const float4*>(xp + k * a.sR_x) : make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.pre_mul != 0.f) {
#pragma unroll
      for (int k = 0; k < RCACHE; ++k) {
        v[k].x = __fmul_rn(v[k].x, a.pre_mul); v[k].y = __fmul_rn(v[k].y, a.pre_mul);
        v[k].z = __fmul_rn(v[k].z, a.pre_mul); v[k].w = __fmul_rn(v[k].w, a.pre_mul);
      }
    } else if (a.pre_div != 0.f) {
#pragma unroll
      for (int k = 0; k < RCACHE; ++k) {
        v[k].x = __fdiv_rn(v[k].x, a.pre_div); v[k].y = __fdiv_rn(v[k].y, a.pre_div);
        v[k].z = __fdiv_rn(v[k].z, a.pre_div); v[k].w = __fdiv_rn(v[k].w, a.pre_div);
      }
    }
    if (a.avoid_overflow) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int k = 0; k < RCACHE; ++k)
        if (k < a.R) { s.x = __fadd_rn(s.x, v[k].x); s.y = __fadd_rn(s.y, v[k].y); s.z = __fadd_rn(s.z, v[k].z); s.w = __fadd_rn(s.w, v[k].w); }
      const float rf = (float)a.R;
      const float4 mean = make_float4(__fdiv_rn(s.x, rf), __fdiv_rn(s.y, rf), __fdiv_rn(s.z, rf), __fdiv_rn(s.w, rf));
#pragma unroll
      for (int k = 0; k < RCACHE; ++k) {
        v[k].x = __fsub_rn(v[k].x, mean.x); v[k].y = __fsub_rn(v[k].y, mean.y);
        v[k].z = __fsub_rn(v[k].z, mean.z); v[k].w = __fsub_rn(v[k].w, mean.w);
      }
    }
    float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < RCACHE; ++k)
      if (k < a.R) {
        v[k].x = expf(v[k].x); v[k].y = expf(v[k].y); v[k].z = expf(v[k].z); v[k].w = expf(v[k].w);
        z.x = __fadd_rn(z.x, v[k].x); z.y = __fadd_rn(z.y, v[k].y); z.z = __fadd_rn(z.z, v[k].z); z.w = __fadd_rn(z.w, v[k].w);
      }
#pragma unroll
    for (int k = 0; k < RCACHE; ++k)
      if (k < a.R)
        *reinterpret_cast<float4*>(op + k * a.sR_o) =
            make_float4(__fdiv_rn(v[k].x, z.x), __fdiv_rn(v[k].y, z.y), __fdiv_rn(v[k].z, z.z), __fdiv_rn(v[k].w, z.w));
  }
}

static inline bool al16s(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ------------------------------------------------------------------ softmax-cross-entropy, summed
// `(!sum (softmax-cross-entropy x labels))` as the MNIST example composes its loss
// (examples/mnist/mlp.lisp:24-28; nn/criterion.lisp:89-117,164-188): z = softmax(x, avoid-overflow nil),
// out = -1 * (labels * log(z + delta)), loss = sum(out).  Three node launches (softmax, the fused
// log/mul chain, `!sum`) become one kernel: a warp per row keeps the row in registers, the row
// losses meet through a last-block ticket in a fixed order.  z is written because the backward
// (`sce-diff-static`, :81-87) needs it.
struct SceArgs {
  const float* x; const float* labels; float* z; float* loss; float* partial; unsigned int* counter;
  int64_t rows, cols;
  float delta;
  const float* dy; float* dx;   // both or neither: dx = (dy[0] * (z - labels)) / rows written as well (`sce-diff-static` in the same launch)
};

__global__ void __launch_bounds__(256) sce_sum_forward_kernel(const __grid_constant__ SceArgs a) {
  __shared__ float smem[8];
  __shared__ bool is_last;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  float rl = 0.f;
  if (row < a.rows) {
    const float* xp = a.x + row * a.cols;
    const float* lp = a.labels + row * a.cols;
    float* zp = a.z + row * a.cols;
    float v[32];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      const int64_t c = lane + 32 * k;
      v[k] = 0.f;
      if (c < a.cols) { v[k] = expf(xp[c]); s = __fadd_rn(s, v[k]); }
    }
    s = warp_sum(s);
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      const int64_t c = lane + 32 * k;
      if (c < a.cols) {
        const float z = __fdiv_rn(v[k], s);
        zp[c] = z;
        if (a.dx) a.dx[row * a.cols + c] = __fdiv_rn(__fmul_rn(*a.dy, __fsub_rn(z, lp[c])), (float)a.rows);   // sce_backward_kernel's operations
        rl = __fadd_rn(rl, __fmul_rn(-1.0f, __fmul_rn(lp[c], logf(__fadd_rn(z, a.delta)))));
      }
    }
  }
  rl = warp_sum(rl);
  if (lane == 0) smem[threadIdx.x >> 5] = rl;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = smem[0];
#pragma unroll
    for (int k = 1; k < 8; ++k) t = __fadd_rn(t, smem[k]);
    unsigned int tk;
    asm volatile("st.relaxed.gpu.global.f32 [%0], %1;" ::"l"(a.partial + blockIdx.x), "f"(t) : "memory");
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], 1;" : "=r"(tk) : "l"(a.counter) : "memory");
    is_last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  float w = 0.f;
  for (uint32_t c = threadIdx.x; c < gridDim.x; c += 256) w = __fadd_rn(w, __ldcg(a.partial + c));
  w = block_sum<256>(w, smem);
  if (threadIdx.x == 0) { a.loss[0] = w; *a.counter = 0; }
}

// `sce-diff-static` (criterion.lisp:81-87): dx = (dy * (z - labels)) / batch, dy the (broadcast) scalar
// gradient of the summed loss, read from device memory.
__global__ void __launch_bounds__(256) sce_backward_kernel(const float* __restrict__ z, const float* __restrict__ labels,
                                                           const float* __restrict__ dy, float* __restrict__ dx, int64_t n,
                                                           float batch, int accumulate) {
  const float d = *dy;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    float g = __fdiv_rn(__fmul_rn(d, __fsub_rn(z[i], labels[i])), batch);
    if (accumulate) g = __fadd_rn(dx[i], g);
    dx[i] = g;
  }
}

}  // namespace wf2

using namespace wf2;

extern "C" int wf2cu_softmax_ex(const wf2cu_view* x, const wf2cu_view* out, int axis, int avoid_overflow, float pre_divisor);
extern "C" int wf2cu_softmax(const wf2cu_view* x, const wf2cu_view* out, int axis, int avoid_overflow) {
  return wf2cu_softmax_ex(x, out, axis, avoid_overflow, 0.f);
}

extern "C" int wf2cu_softmax_ex(const wf2cu_view* x, const wf2cu_view* out, int axis, int avoid_overflow, float pre_divisor) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(x && out && x->base && out->base, "wf2cu_softmax: null view");
  WF2_CHECK(x->rank == out->rank && x->rank >= 1 && x->rank <= kMaxRank, "wf2cu_softmax: bad rank");
  const int rank = x->rank;
  if (axis < 0) axis += rank;
  WF2_CHECK(axis >= 0 && axis < rank, "wf2cu_softmax: axis %d out of range for rank %d", axis, rank);
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) {
    WF2_CHECK(x->shape[i] == out->shape[i], "wf2cu_softmax: shape mismatch on axis %d", i);
    numel *= x->shape[i];
  }
  if (numel == 0) return 0;
  State& s = S();
  const float* xp = fptr(x);
  float* op = fptr(out);
  const int64_t R = x->shape[axis];
  // is the softmax axis the innermost non-trivial one?
  bool innermost = true;
  for (int i = axis + 1; i < rank; ++i) if (x->shape[i] != 1) innermost = false;
  if (innermost && pre_divisor != 0.f)
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_softmax_ex: the folded divisor is implemented for a non-innermost softmax axis "
                     "only; issue the !div node and wf2cu_softmax separately");
  if (innermost) {
    SmRowArgs a;
    memset(&a, 0, sizeof(a));
    a.x = xp; a.out = op; a.cols = R; a.avoid_overflow = avoid_overflow;
    // merge outer axes
    int n = 0;
    int64_t rows = 1;
    bool ok = true;
    bool strides4 = true;
    for (int i = 0; i < axis; ++i) {
      const int64_t e = x->shape[i];
      if (e == 1) continue;
      rows *= e;
      if (n > 0 && a.ostride_x[n - 1] == x->stride[i] * e && a.ostride_o[n - 1] == out->stride[i] * e) {
        a.oshape[n - 1] *= (uint32_t)e; a.ostride_x[n - 1] = x->stride[i]; a.ostride_o[n - 1] = out->stride[i];
      } else {
        if (n == 4) { ok = false; break; }
        a.oshape[n] = (uint32_t)e; a.ostride_x[n] = x->stride[i]; a.ostride_o[n] = out->stride[i]; ++n;
      }
    }
    for (int i = 0; i < n; ++i) if (a.ostride_x[i] % 4 || a.ostride_o[i] % 4) strides4 = false;
    if (ok && rows < (int64_t(1) << 31)) {
      a.nouter = n; a.rows = (uint32_t)rows;
      const int64_t sx = x->stride[axis], so = out->stride[axis];
      const bool vec = (sx == 1 && so == 1 && R % 4 == 0 && al16s(xp) && al16s(op) && strides4);
      SmWarpExtra e{sx, so};
      if (R <= 1024 && !(vec && R > 256)) {
        softmax_row_warp_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s.stream>>>(a, e);
        return post_launch("wf2cu_softmax(row,warp)");
      }
      if (vec && R <= 1024) {
        WF2_CUDA(launch_pdl(softmax_row_block_kernel<128, 2>, dim3((unsigned)rows), dim3(128), 0, s.stream, a));
        return post_launch("wf2cu_softmax(row,block128)");
      }
      if (vec && R <= 4096) {
        WF2_CUDA(launch_pdl(softmax_row_block_kernel<128, 8>, dim3((unsigned)rows), dim3(128), 0, s.stream, a));
        return post_launch("wf2cu_softmax(row,block128x8)");
      }
      if (vec && R <= 16384) {
        WF2_CUDA(launch_pdl(softmax_row_block_kernel<512, 8>, dim3((unsigned)rows), dim3(512), 0, s.stream, a));
        return post_launch("wf2cu_softmax(row,block512)");
      }
      softmax_row_stream_kernel<<<(unsigned)rows, 512, 0, s.stream>>>(a, e);
      return post_launch("wf2cu_softmax(row,stream)");
    }
  }
  // lane kernel
  SmLaneArgs a;
  memset(&a, 0, sizeof(a));
  a.x = xp; a.out = op; a.R = R; a.sR_x = x->stride[axis]; a.sR_o = out->stride[axis];
  a.avoid_overflow = avoid_overflow;
  a.pre_div = pre_divisor;
  {
    int e = 0;
    const float mant = frexpf(fabsf(pre_divisor), &e);
    // power of two with a representable, normal reciprocal
    if (pre_divisor != 0.f && mant == 0.5f && e > -120 && e < 120) a.pre_mul = 1.0f / pre_divisor;
  }
  a.nlanes = 1;
  for (int i = 0; i < rank; ++i) {
    if (i == axis || x->shape[i] == 1) continue;
    a.lshape[a.nlane] = x->shape[i]; a.lstride_x[a.nlane] = x->stride[i]; a.lstride_o[a.nlane] = out->stride[i];
    a.nlanes *= x->shape[i]; ++a.nlane;
  }
  // 4 lanes per thread when the innermost lane axis is contiguous on both sides and everything is 16-byte aligned
  if (R <= 16 && a.nlane >= 1 && a.lstride_x[a.nlane - 1] == 1 && a.lstride_o[a.nlane - 1] == 1 && a.lshape[a.nlane - 1] % 4 == 0 &&
      al16s(xp) && al16s(op) && a.sR_x % 4 == 0 && a.sR_o % 4 == 0 && a.nlanes >= (int64_t)s.sm_count * 1024) {
    bool ok = true;
    for (int i = 0; i + 1 < a.nlane; ++i) ok = ok && a.lstride_x[i] % 4 == 0 && a.lstride_o[i] % 4 == 0;
    if (ok) {
      SmLaneArgs v = a;
      v.lshape[v.nlane - 1] /= 4; v.lstride_x[v.nlane - 1] = 4; v.lstride_o[v.nlane - 1] = 4;
      v.nlanes /= 4;
      int64_t g4 = (v.nlanes + 255) / 256;
      const int64_t cap4 = (int64_t)s.sm_count * 16;
      if (g4 > cap4) g4 = cap4;
      if (R <= 8) WF2_CUDA(launch_pdl(softmax_lane_vec_kernel<8>, dim3((unsigned)g4), dim3(256), 0, s.stream, v));
      else if (R <= 12) WF2_CUDA(launch_pdl(softmax_lane_vec_kernel<12>, dim3((unsigned)g4), dim3(256), 0, s.stream, v));
      else WF2_CUDA(launch_pdl(softmax_lane_vec_kernel<16>, dim3((unsigned)g4), dim3(256), 0, s.stream, v));
      return post_launch("wf2cu_softmax(lane,vec4)");
    }
  }
  int64_t grid = (a.nlanes + 255) / 256;
  const int64_t cap = (int64_t)s.sm_count * 16;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  if (R <= 16) softmax_lane_kernel<16><<<(unsigned)grid, 256, 0, s.stream>>>(a);
  else if (R <= 64) softmax_lane_kernel<64><<<(unsigned)grid, 256, 0, s.stream>>>(a);
  else softmax_lane_kernel<0><<<(unsigned)grid, 256, 0, s.stream>>>(a);
  return post_launch("wf2cu_softmax(lane)");
}

// ------------------------------------------------------------------ LayerNorm (GPT-2 example)
// `LayerNorm-Revisit` (examples/gpt-2/utils.lisp:114-127) is ten nodes: !mean, !sub, !expt 2, !mean, !add eps,
// !sqrt, !div, !mul g, !add b (+ the broadcast views) — ten passes over a (B N E) tensor, 225 launches of the 562 of
// a GPT-2 small forward.  One kernel, a warp per row with the row in registers; every element goes through the
// same operations in the same order as the node chain (mean = sum / count with a correctly rounded divide, square
// as a product, IEEE sqrt and divide), the two row sums are shuffle trees (the 1e-5 reduction bar).
struct LnArgs { const float* x; const float* g; const float* b; float* out; int64_t rows; int cols; float eps;
                float* plane; };   // != NULL: also the mixed bf16 operand plane of the result (planes_emit_begin)

template <int V>   // float4 per lane: cols <= 128 * V
__global__ void __launch_bounds__(256) layernorm_kernel(const __grid_constant__ LnArgs a) {
  pdl_prologue();
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= a.rows) return;
  const int nv = a.cols >> 2;
  const float4* xr = reinterpret_cast<const float4*>(a.x + row * a.cols);
  float4 v[V];
#pragma unroll
  for (int k = 0; k < V; ++k) { const int i = lane + 32 * k; v[k] = i < nv ? xr[i] : make_float4(0.f, 0.f, 0.f, 0.f); }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < V; ++k) s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float cnt = (float)a.cols;
  const float u = __fdiv_rn(s, cnt);
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int i = lane + 32 * k;
    v[k].x = __fsub_rn(v[k].x, u); v[k].y = __fsub_rn(v[k].y, u); v[k].z = __fsub_rn(v[k].z, u); v[k].w = __fsub_rn(v[k].w, u);
    if (i < nv) q += (__fmul_rn(v[k].x, v[k].x) + __fmul_rn(v[k].y, v[k].y)) + (__fmul_rn(v[k].z, v[k].z) + __fmul_rn(v[k].w, v[k].w));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float r = __fsqrt_rn(__fadd_rn(__fdiv_rn(q, cnt), a.eps));
  const float4* gr = reinterpret_cast<const float4*>(a.g);
  const float4* br = reinterpret_cast<const float4*>(a.b);
  float4* orow = reinterpret_cast<float4*>(a.out + row * a.cols);
#pragma unroll
  for (int k = 0; k < V; ++k) {
    const int i = lane + 32 * k;
    if (i < nv) {
      const float4 g4 = __ldg(gr + i), b4 = __ldg(br + i);
      float4 y;
      y.x = __fadd_rn(__fmul_rn(__fdiv_rn(v[k].x, r), g4.x), b4.x);
      y.y = __fadd_rn(__fmul_rn(__fdiv_rn(v[k].y, r), g4.y), b4.y);
      y.z = __fadd_rn(__fmul_rn(__fdiv_rn(v[k].z, r), g4.z), b4.z);
      y.w = __fadd_rn(__fmul_rn(__fdiv_rn(v[k].w, r), g4.w), b4.w);
      orow[i] = y;
      if (a.plane) plane_store4(a.plane + row * a.cols, 4 * i, y.x, y.y, y.z, y.w);
    }
  }
}

extern "C" int wf2cu_layernorm(const void* x, int64_t x_off, const void* g, int64_t g_off, const void* b, int64_t b_off,
                               void* out, int64_t out_off, int64_t rows, int64_t cols, float eps) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(x && g && b && out && rows >= 0 && cols >= 1, "wf2cu_layernorm: bad arguments");
  LnArgs a{(const float*)x + x_off, (const float*)g + g_off, (const float*)b + b_off, (float*)out + out_off, rows, (int)cols, eps, nullptr};
  const bool ok = cols % 4 == 0 && cols <= 1024 && al16s(a.x) && al16s(a.g) && al16s(a.b) && al16s(a.out);
  if (!ok)
    return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_layernorm: rows of %lld floats (need a multiple of 4, <= 1024, 16-byte aligned): "
                     "issue the node chain instead", (long long)cols);
  if (rows == 0) return 0;
  const int64_t grid = (rows + 7) / 8;
  WF2_CHECK(grid < (int64_t(1) << 31), "wf2cu_layernorm: too many rows");
  if (S().emit_next) a.plane = planes_emit_begin(rows, cols);      // wf2cu_emit_operand_planes: one-shot, may decline
  cudaStream_t st = S().stream;
  if (cols <= 256) WF2_CUDA(launch_pdl(layernorm_kernel<2>, dim3((unsigned)grid), dim3(256), 0, st, a));
  else if (cols <= 512) WF2_CUDA(launch_pdl(layernorm_kernel<4>, dim3((unsigned)grid), dim3(256), 0, st, a));
  else WF2_CUDA(launch_pdl(layernorm_kernel<8>, dim3((unsigned)grid), dim3(256), 0, st, a));
  const int rc = post_launch("wf2cu_layernorm");
  if (!rc && a.plane) planes_emit_publish(a.out, rows, cols, cols, a.plane);
  return rc;
}

extern "C" int wf2cu_sce_sum_forward_grad(const void* x, int64_t x_off, const void* labels, int64_t l_off, void* z, int64_t z_off,
                                          void* loss, int64_t loss_off, int64_t rows, int64_t cols, float delta,
                                          const void* dy, int64_t dy_off, void* dx, int64_t dx_off);
extern "C" int wf2cu_sce_sum_forward(const void* x, int64_t x_off, const void* labels, int64_t l_off, void* z, int64_t z_off,
                                     void* loss, int64_t loss_off, int64_t rows, int64_t cols, float delta) {
  return wf2cu_sce_sum_forward_grad(x, x_off, labels, l_off, z, z_off, loss, loss_off, rows, cols, delta, nullptr, 0, nullptr, 0);
}

extern "C" int wf2cu_sce_sum_forward_grad(const void* x, int64_t x_off, const void* labels, int64_t l_off, void* z, int64_t z_off,
                                          void* loss, int64_t loss_off, int64_t rows, int64_t cols, float delta,
                                          const void* dy, int64_t dy_off, void* dx, int64_t dx_off) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(x && labels && z && loss && rows >= 0 && cols >= 1, "wf2cu_sce_sum_forward: bad arguments");
  WF2_CHECK((dy != nullptr) == (dx != nullptr), "wf2cu_sce_sum_forward_grad: dy and dx come together");
  if (cols > 1024) return set_error(WF2CU_ERR_UNSUPPORTED, "wf2cu_sce_sum_forward: %lld classes > 1024; compose softmax, the log chain and !sum instead", (long long)cols);
  const int64_t grid = rows == 0 ? 1 : (rows + 7) / 8;
  WF2_CHECK(grid < (int64_t(1) << 31), "wf2cu_sce_sum_forward: too many rows");
  int rc = ensure_scratch((size_t)grid * sizeof(float));
  if (rc) return rc;
  State& s = S();
  SceArgs a{(const float*)x + x_off, (const float*)labels + l_off, (float*)z + z_off, (float*)loss + loss_off, (float*)scratch_base(),
            counters_base() + 4095, rows, cols, delta, dy ? (const float*)dy + dy_off : nullptr, dx ? (float*)dx + dx_off : nullptr};
  sce_sum_forward_kernel<<<(unsigned)grid, 256, 0, s.stream>>>(a);
  return post_launch("wf2cu_sce_sum_forward");
}

extern "C" int wf2cu_sce_backward(const void* z, int64_t z_off, const void* labels, int64_t l_off, const void* dy, int64_t dy_off,
                                  void* dx, int64_t dx_off, int64_t rows, int64_t cols, int accumulate) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(z && labels && dy && dx && rows >= 0 && cols >= 0, "wf2cu_sce_backward: bad arguments");
  const int64_t n = rows * cols;
  if (n == 0) return 0;
  int64_t grid = (n + 255) / 256;
  const int64_t cap = (int64_t)S().sm_count * 8;
  if (grid > cap) grid = cap;
  sce_backward_kernel<<<(unsigned)grid, 256, 0, S().stream>>>((const float*)z + z_off, (const float*)labels + l_off,
                                                             (const float*)dy + dy_off, (float*)dx + dx_off, n, (float)rows, accumulate);
  return post_launch("wf2cu_sce_backward");
}
