// conv.cu — unfold (Im2Col / Col2Im), the fused unfold/fold pair Conv2D uses, and 2-D pooling
// (SURVEY §8 f-3).  Replaces `waffe2_im2col_s_column` / `waffe2_col2im_s_column`
// (`cl-waffe2-simd/kernels/im2col.c:63-115,395-448`, bound in `cl-waffe2-simd/api.lisp:27-61`, called
// from `source/backends/cpu/im2col.lisp:4-50`) and the node chains `nn/im2col.lisp:15-98`,
// `nn/conv.lisp:88-105`, `nn/pool.lisp:73-105` build around them.
//
// Everything here is data movement plus short fixed-order sums: HBM/L2-bound, one thread per output
// element with the fastest-varying thread index on the contiguous axis of the side that moves more
// bytes.  Sums run in the reference's order (kernel offsets (kh, kw) ascending, starting from 0.0f) so
// results are bit-identical to the C kernels.
//
// Reference semantics kept as written (im2col.c:41,43): the node sees an already padded image and visits
// image rows y = kh + j*stride_h only while y < dilation_h*kh - pad_h + stride_h*H_out, i.e. (dilation 1)
// while j*stride_h + pad_h < stride_h*H_out — with pad > 0 the last output rows/columns are never
// touched.  Dilation > 1 walks out of the buffer in the reference ("currently not working, please set 1",
// nn/conv.lisp:41) and is declined here.
#include "common.cuh"

namespace wf2 {

struct Geom {
  int N, C, H, W;          // image extents as the kernel indexes them
  int Ho, Wo, KH, KW;
  int pad_h, pad_w, sh, sw;
  int jmax, imax;          // number of output rows / columns the reference visits
};

static int make_geom(const wf2cu_unfold_geom* g, Geom* o, const char* who) {
  WF2_CHECK(g != nullptr, "%s: null geometry", who);
  WF2_CHECK(g->N >= 0 && g->C >= 0 && g->H >= 0 && g->W >= 0 && g->Ho >= 0 && g->Wo >= 0, "%s: negative extent", who);
  WF2_CHECK(g->KH >= 1 && g->KW >= 1 && g->stride_h >= 1 && g->stride_w >= 1 && g->pad_h >= 0 && g->pad_w >= 0,
            "%s: kernel / stride must be >= 1 and padding >= 0", who);
  if (g->dilation_h != 1 || g->dilation_w != 1)
    return set_error(WF2CU_ERR_UNSUPPORTED, "%s: dilation (%lld %lld) != 1 — the reference's unfold walks out of its buffer "
                     "for dilation > 1 (nn/conv.lisp:41)", who, (long long)g->dilation_h, (long long)g->dilation_w);
  const int64_t lim = int64_t(1) << 30;
  WF2_CHECK(g->N < lim && g->C < lim && g->H < lim && g->W < lim && g->Ho < lim && g->Wo < lim && g->KH < lim && g->KW < lim,
            "%s: extent too large", who);
  o->N = (int)g->N; o->C = (int)g->C; o->H = (int)g->H; o->W = (int)g->W;
  o->Ho = (int)g->Ho; o->Wo = (int)g->Wo; o->KH = (int)g->KH; o->KW = (int)g->KW;
  o->pad_h = (int)g->pad_h; o->pad_w = (int)g->pad_w; o->sh = (int)g->stride_h; o->sw = (int)g->stride_w;
  // j is visited iff j*sh + pad_h < sh*Ho  <=>  j < ceil((sh*Ho - pad_h) / sh)
  const int64_t eh = g->stride_h * g->Ho - g->pad_h, ew = g->stride_w * g->Wo - g->pad_w;
  o->jmax = eh <= 0 ? 0 : (int)((eh + g->stride_h - 1) / g->stride_h);
  o->imax = ew <= 0 ? 0 : (int)((ew + g->stride_w - 1) / g->stride_w);
  return 0;
}

static inline unsigned grid_for(int64_t n) {
  int64_t g = (n + 255) / 256;
  const int64_t cap = (int64_t)S().sm_count * 32;
  if (g > cap) g = cap;
  return (unsigned)(g < 1 ? 1 : g);
}

// ------------------------------------------------------------------ Im2ColNode / Col2ImNode
// col is (N C KH KW Ho Wo) contiguous, im is (N C H W) contiguous (already padded).
__global__ void __launch_bounds__(256) im2col_kernel(float* __restrict__ col, const float* __restrict__ im, const Geom g, int64_t total) {
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    int64_t r = t;
    const int i = (int)(r % g.Wo); r /= g.Wo;
    const int j = (int)(r % g.Ho); r /= g.Ho;
    const int kw = (int)(r % g.KW); r /= g.KW;
    const int kh = (int)(r % g.KH); r /= g.KH;          // r = n*C + c
    if (j < g.jmax && i < g.imax)
      col[t] = im[(r * g.H + (kh + j * g.sh)) * g.W + (kw + i * g.sw)];
  }
}

__global__ void __launch_bounds__(256) col2im_kernel(const float* __restrict__ col, float* __restrict__ im, const Geom g, int64_t total) {
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    int64_t r = t;
    const int x = (int)(r % g.W); r /= g.W;
    const int y = (int)(r % g.H); r /= g.H;             // r = n*C + c
    float acc = 0.0f;
    for (int kh = 0; kh < g.KH && kh <= y; ++kh) {
      const int dy = y - kh;
      if (dy % g.sh) continue;
      const int j = dy / g.sh;
      if (j >= g.jmax) continue;
      for (int kw = 0; kw < g.KW && kw <= x; ++kw) {
        const int dx = x - kw;
        if (dx % g.sw) continue;
        const int i = dx / g.sw;
        if (i >= g.imax) continue;
        acc = __fadd_rn(acc, col[(((r * g.KH + kh) * g.KW + kw) * g.Ho + j) * g.Wo + i]);
      }
    }
    im[t] = acc;
  }
}

// ------------------------------------------------------------------ fused unfold / fold
// unfold: x is the UNPADDED (N C Hin Win) input as a strided view; rows is (N*Ho*Wo, C*KH*KW) contiguous.
//   rows[(n j i), (c kh kw)] = padded(x)[n, c, kh + j*sh, kw + i*sw]   where the node visits (j, i), else 0
// = padding + Im2ColNode + permute (0 4 5 1 2 3) + ->contiguous + reshape in one pass.
struct View4 { const float* p; int64_t s[4]; };

__device__ __forceinline__ float sample(const View4& x, const Geom& g, int n, int c, int yp, int xp) {
  const int y = yp - g.pad_h, xx = xp - g.pad_w;
  if (y < 0 || y >= g.H || xx < 0 || xx >= g.W) return 0.0f;
  return x.p[n * x.s[0] + c * x.s[1] + y * x.s[2] + xx * x.s[3]];
}

__global__ void __launch_bounds__(256) unfold_rows_kernel(float* __restrict__ rows, const View4 x, const Geom g, int64_t total) {
  // one thread per (n, j, i, c, kh): the index arithmetic is paid once per KW outputs
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    int64_t r = t;
    const int kh = (int)(r % g.KH); r /= g.KH;
    const int c = (int)(r % g.C); r /= g.C;
    const int i = (int)(r % g.Wo); r /= g.Wo;
    const int j = (int)(r % g.Ho); r /= g.Ho;
    const int n = (int)r;
    float* o = rows + t * g.KW;
    const int y = kh + j * g.sh - g.pad_h;
    const bool row_ok = j < g.jmax && i < g.imax && y >= 0 && y < g.H;
    const float* src = x.p + n * x.s[0] + c * x.s[1] + (int64_t)y * x.s[2];
    const int x0 = i * g.sw - g.pad_w;
    for (int kw = 0; kw < g.KW; ++kw) {
      const int xx = x0 + kw;
      o[kw] = (row_ok && xx >= 0 && xx < g.W) ? src[xx * x.s[3]] : 0.0f;
    }
  }
}

// fold: dx[n, c, y, x] (+)= sum over (kh, kw) ascending of drows[(n j i), (c kh kw)] for the windows that
// visited padded position (y + pad_h, x + pad_w)
// = reshape + permute back + Col2ImNode + the padding window + the gradient adder in one pass.
struct View4W { float* p; int64_t s[4]; };

__global__ void __launch_bounds__(256) fold_rows_kernel(const float* __restrict__ drows, const View4W dx, const Geom g, int64_t total, int accumulate) {
  const int K = g.C * g.KH * g.KW;
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    int64_t r = t;
    const int xi = (int)(r % g.W); r /= g.W;
    const int yi = (int)(r % g.H); r /= g.H;
    const int c = (int)(r % g.C); r /= g.C;
    const int n = (int)r;
    const int y = yi + g.pad_h, x = xi + g.pad_w;
    float acc = 0.0f;
    for (int kh = 0; kh < g.KH && kh <= y; ++kh) {
      const int dy = y - kh;
      if (dy % g.sh) continue;
      const int j = dy / g.sh;
      if (j >= g.jmax) continue;
      for (int kw = 0; kw < g.KW && kw <= x; ++kw) {
        const int dxx = x - kw;
        if (dxx % g.sw) continue;
        const int i = dxx / g.sw;
        if (i >= g.imax) continue;
        acc = __fadd_rn(acc, drows[((int64_t)(n * g.Ho + j) * g.Wo + i) * K + (c * g.KH + kh) * g.KW + kw]);
      }
    }
    float* q = dx.p + n * dx.s[0] + c * dx.s[1] + yi * dx.s[2] + xi * dx.s[3];
    *q = accumulate ? __fadd_rn(*q, acc) : acc;
  }
}

// Same result, staged: for one (n, c) the contributions live in Ho*Wo segments of KH*KW consecutive floats
// (one per output position).  A CTA copies that slab into shared memory with coalesced reads and builds the
// (H, W) plane from it, so every float of drows is read from global memory once instead of a 32-byte sector per
// contribution (553 MB of sector traffic for 33 MB of data on MNIST-CNN's second convolution at batch 1024:
// 54.6 us).  Used when the slab fits 40 KiB.  Sum order unchanged.
__global__ void __launch_bounds__(256) fold_rows_slab_kernel(const float* __restrict__ drows, const View4W dx, const Geom g, int accumulate) {
  extern __shared__ float slab[];                        // [Ho*Wo][KH*KW]
  const int c = blockIdx.x % g.C, n = blockIdx.x / g.C;
  const int KK = g.KH * g.KW, K = g.C * KK, P = g.Ho * g.Wo;
  const float* base = drows + (int64_t)n * P * K + c * KK;
  {
    // 8 unconditional (clamped) loads per thread in flight: the copy is a chain of round trips otherwise
    const int total = P * KK;
    for (int e0 = threadIdx.x; e0 < total; e0 += 256 * 8) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int e = min(e0 + u * 256, total - 1);
        const int pos = e / KK, q = e - pos * KK;
        v[u] = __ldg(base + (int64_t)pos * K + q);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) { const int e = e0 + u * 256; if (e < total) slab[e] = v[u]; }
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < g.H * g.W; e += 256) {
    const int yi = e / g.W, xi = e - yi * g.W;
    const int y = yi + g.pad_h, x = xi + g.pad_w;
    float acc = 0.0f;
    for (int kh = 0; kh < g.KH && kh <= y; ++kh) {
      const int dy = y - kh;
      if (dy % g.sh) continue;
      const int j = dy / g.sh;
      if (j >= g.jmax) continue;
      for (int kw = 0; kw < g.KW && kw <= x; ++kw) {
        const int dxx = x - kw;
        if (dxx % g.sw) continue;
        const int i = dxx / g.sw;
        if (i >= g.imax) continue;
        acc = __fadd_rn(acc, slab[(j * g.Wo + i) * KK + kh * g.KW + kw]);
      }
    }
    float* q = dx.p + n * dx.s[0] + c * dx.s[1] + yi * dx.s[2] + xi * dx.s[3];
    *q = accumulate ? __fadd_rn(*q, acc) : acc;
  }
}

// ------------------------------------------------------------------ pooling
// MaxPool2D / AvgPool2D (nn/pool.lisp:73-105) = unfold -> (.., KH*KW) windows -> !max / !mean over axis 1
// -> (N Ho Wo C) -> permute to (N C Ho Wo).  One pass: each thread owns one (n, c, j, i) window, walks it
// in (kh, kw) order (the order the window's row has in memory in the reference, which is the order its
// sequential fp32 sum runs in) and writes through the output view.
// C_FAST: consecutive threads take consecutive channels (input stored channels-last, as Conv2D leaves
// it); otherwise consecutive threads take consecutive output columns (plain NCHW input).
template <bool AVG, bool C_FAST>
__global__ void __launch_bounds__(256) pool2d_kernel(const View4 x, const View4W out, const Geom g, int64_t total) {
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    int64_t r = t;
    int n, c, j, i;
    if (C_FAST) { c = (int)(r % g.C); r /= g.C; i = (int)(r % g.Wo); r /= g.Wo; j = (int)(r % g.Ho); r /= g.Ho; n = (int)r; }
    else        { i = (int)(r % g.Wo); r /= g.Wo; j = (int)(r % g.Ho); r /= g.Ho; c = (int)(r % g.C); r /= g.C; n = (int)r; }
    const bool visited = j < g.jmax && i < g.imax;
    float acc = 0.0f;
    for (int kh = 0; kh < g.KH; ++kh)
      for (int kw = 0; kw < g.KW; ++kw) {
        const float v = visited ? sample(x, g, n, c, kh + j * g.sh, kw + i * g.sw) : 0.0f;
        if (AVG) acc = __fadd_rn(acc, v);
        else acc = (kh == 0 && kw == 0) ? v : fmaxf(acc, v);
      }
    if (AVG) acc = __fdiv_rn(acc, (float)(g.KH * g.KW));
    out.p[n * out.s[0] + c * out.s[1] + j * out.s[2] + i * out.s[3]] = acc;
  }
}

// backward.  max (MaxValue-Node backward, base-impl/matrix-ops.lisp:311-334): mask = (a == max a),
// grad = mask * dout, then the unfold backward; avg (`!mean` = !sum then !div): dout / count spread over
// the window, then the unfold backward.  Thread per input element, windows in (kh, kw) order.
template <bool AVG, bool C_FAST>
__global__ void __launch_bounds__(256) pool2d_backward_kernel(const View4 x, const View4 out, const View4 dout, const View4W dx,
                                                              const Geom g, int64_t total, int accumulate, int relu_input) {
  const float cnt = (float)(g.KH * g.KW);
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (int64_t)gridDim.x * 256) {
    int64_t r = t;
    int n, c, yi, xi;
    if (C_FAST) { c = (int)(r % g.C); r /= g.C; xi = (int)(r % g.W); r /= g.W; yi = (int)(r % g.H); r /= g.H; n = (int)r; }
    else        { xi = (int)(r % g.W); r /= g.W; yi = (int)(r % g.H); r /= g.H; c = (int)(r % g.C); r /= g.C; n = (int)r; }
    const int y = yi + g.pad_h, xx = xi + g.pad_w;
    const float v = (AVG && !relu_input) ? 0.0f : x.p[n * x.s[0] + c * x.s[1] + yi * x.s[2] + xi * x.s[3]];
    float acc = 0.0f;
    for (int kh = 0; kh < g.KH && kh <= y; ++kh) {
      const int dy = y - kh;
      if (dy % g.sh) continue;
      const int j = dy / g.sh;
      if (j >= g.jmax) continue;
      for (int kw = 0; kw < g.KW && kw <= xx; ++kw) {
        const int dxx = xx - kw;
        if (dxx % g.sw) continue;
        const int i = dxx / g.sw;
        if (i >= g.imax) continue;
        const int64_t oo = n * dout.s[0] + c * dout.s[1] + j * dout.s[2] + i * dout.s[3];
        float gcontrib;
        if (AVG) gcontrib = __fdiv_rn(dout.p[oo], cnt);
        else {
          const float m = out.p[n * out.s[0] + c * out.s[1] + j * out.s[2] + i * out.s[3]];
          gcontrib = __fmul_rn(v == m ? 1.0f : 0.0f, dout.p[oo]);
        }
        acc = __fadd_rn(acc, gcontrib);
      }
    }
    // x is the output of a fused `... + !relu` node: its incoming gradient passes the MulNode backward of
    // `!relu`, dout * (x > 0 ? 1 : 0) (x > 0 <=> pre-activation > 0)
    if (relu_input) acc = __fmul_rn(acc, v > 0.f ? 1.f : 0.f);
    float* q = dx.p + n * dx.s[0] + c * dx.s[1] + yi * dx.s[2] + xi * dx.s[3];
    *q = accumulate ? __fadd_rn(*q, acc) : acc;
  }
}

// Channels-last variant (every view has channel stride 1 — what Conv2D and the pooling layers hand over): one
// thread per pixel (n, y, x) finds its windows once and then walks the C channels, which are consecutive in
// memory.  The index arithmetic (several integer divisions) is paid once per pixel instead of once per element:
// 54 -> see profiles on (1024 4 26 26).  At most 16 windows per pixel (kernel / stride ratio), else the kernel above.
template <bool AVG>
__global__ void __launch_bounds__(256) pool2d_backward_cl_kernel(const View4 x, const View4 out, const View4 dout, const View4W dx,
                                                                 const Geom g, int64_t pixels, int accumulate, int relu_input) {
  const float cnt = (float)(g.KH * g.KW);
  for (int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x; t < pixels; t += (int64_t)gridDim.x * 256) {
    int64_t r = t;
    const int xi = (int)(r % g.W); r /= g.W;
    const int yi = (int)(r % g.H); r /= g.H;
    const int n = (int)r;
    const int y = yi + g.pad_h, xx = xi + g.pad_w;
    int64_t wo[16], wd[16];
    int nw = 0;
    for (int kh = 0; kh < g.KH && kh <= y; ++kh) {
      const int dy = y - kh;
      if (dy % g.sh) continue;
      const int j = dy / g.sh;
      if (j >= g.jmax) continue;
      for (int kw = 0; kw < g.KW && kw <= xx; ++kw) {
        const int dxx = xx - kw;
        if (dxx % g.sw) continue;
        const int i = dxx / g.sw;
        if (i >= g.imax) continue;
        if (nw < 16) {
          wo[nw] = n * out.s[0] + j * out.s[2] + i * out.s[3];
          wd[nw] = n * dout.s[0] + j * dout.s[2] + i * dout.s[3];
        }
        ++nw;
      }
    }
    const float* xp = (AVG && !relu_input) ? nullptr : x.p + n * x.s[0] + yi * x.s[2] + xi * x.s[3];
    float* q = dx.p + n * dx.s[0] + yi * dx.s[2] + xi * dx.s[3];
    for (int c = 0; c < g.C; ++c) {
      const float v = xp ? xp[c] : 0.0f;
      float acc = 0.0f;
      for (int w = 0; w < nw; ++w) {
        float gcontrib;
        if (AVG) gcontrib = __fdiv_rn(dout.p[wd[w] + c], cnt);
        else gcontrib = __fmul_rn(v == out.p[wo[w] + c] ? 1.0f : 0.0f, dout.p[wd[w] + c]);
        acc = __fadd_rn(acc, gcontrib);
      }
      if (relu_input) acc = __fmul_rn(acc, v > 0.f ? 1.f : 0.f);
      q[c] = accumulate ? __fadd_rn(q[c], acc) : acc;
    }
  }
}

static int view4(const wf2cu_view* v, const char* who, int64_t e0, int64_t e1, int64_t e2, int64_t e3, const float** p, int64_t s[4]) {
  WF2_CHECK(v && v->base, "%s: null view", who);
  WF2_CHECK(v->rank == 4, "%s: expected a rank-4 (N C H W) view, got rank %d", who, v->rank);
  WF2_CHECK(v->shape[0] == e0 && v->shape[1] == e1 && v->shape[2] == e2 && v->shape[3] == e3,
            "%s: view shape (%lld %lld %lld %lld) does not match the geometry (%lld %lld %lld %lld)", who,
            (long long)v->shape[0], (long long)v->shape[1], (long long)v->shape[2], (long long)v->shape[3],
            (long long)e0, (long long)e1, (long long)e2, (long long)e3);
  *p = fptr(v);
  for (int d = 0; d < 4; ++d) s[d] = v->stride[d];
  return 0;
}

static int writable4(const wf2cu_view* v, const char* who) {
  for (int d = 0; d < 4; ++d)
    WF2_CHECK(!(v->stride[d] == 0 && v->shape[d] > 1), "%s: destination view has a broadcast axis", who);
  return 0;
}

}  // namespace wf2

using namespace wf2;

extern "C" int wf2cu_im2col(void* col, int64_t col_off, const void* im, int64_t im_off, const wf2cu_unfold_geom* geom) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(col && im, "wf2cu_im2col: null pointer");
  Geom g;
  int rc = make_geom(geom, &g, "wf2cu_im2col");
  if (rc) return rc;
  WF2_CHECK((int64_t)(g.KH - 1) + (int64_t)(g.jmax > 0 ? g.jmax - 1 : 0) * g.sh < (g.H > 0 ? g.H : 1) &&
            (int64_t)(g.KW - 1) + (int64_t)(g.imax > 0 ? g.imax - 1 : 0) * g.sw < (g.W > 0 ? g.W : 1),
            "wf2cu_im2col: the walk leaves the (%d %d) image", g.H, g.W);
  const int64_t total = (int64_t)g.N * g.C * g.KH * g.KW * g.Ho * g.Wo;
  if (total == 0) return 0;
  im2col_kernel<<<grid_for(total), 256, 0, S().stream>>>((float*)col + col_off, (const float*)im + im_off, g, total);
  return post_launch("wf2cu_im2col");
}

extern "C" int wf2cu_col2im(const void* col, int64_t col_off, void* im, int64_t im_off, const wf2cu_unfold_geom* geom) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(col && im, "wf2cu_col2im: null pointer");
  Geom g;
  int rc = make_geom(geom, &g, "wf2cu_col2im");
  if (rc) return rc;
  WF2_CHECK((int64_t)(g.KH - 1) + (int64_t)(g.jmax > 0 ? g.jmax - 1 : 0) * g.sh < (g.H > 0 ? g.H : 1) &&
            (int64_t)(g.KW - 1) + (int64_t)(g.imax > 0 ? g.imax - 1 : 0) * g.sw < (g.W > 0 ? g.W : 1),
            "wf2cu_col2im: the walk leaves the (%d %d) image", g.H, g.W);
  const int64_t total = (int64_t)g.N * g.C * g.H * g.W;
  if (total == 0) return 0;
  col2im_kernel<<<grid_for(total), 256, 0, S().stream>>>((const float*)col + col_off, (float*)im + im_off, g, total);
  return post_launch("wf2cu_col2im");
}

extern "C" int wf2cu_unfold(const wf2cu_view* x, void* rows, int64_t rows_off, const wf2cu_unfold_geom* geom) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(rows, "wf2cu_unfold: null pointer");
  Geom g;
  int rc = make_geom(geom, &g, "wf2cu_unfold");
  if (rc) return rc;
  View4 xv;
  rc = view4(x, "wf2cu_unfold", g.N, g.C, g.H, g.W, &xv.p, xv.s);
  if (rc) return rc;
  const int64_t total = (int64_t)g.N * g.Ho * g.Wo * g.C * g.KH;
  if (total == 0) return 0;
  unfold_rows_kernel<<<grid_for(total), 256, 0, S().stream>>>((float*)rows + rows_off, xv, g, total);
  return post_launch("wf2cu_unfold");
}

extern "C" int wf2cu_fold(const void* drows, int64_t drows_off, const wf2cu_view* dx, const wf2cu_unfold_geom* geom, int accumulate) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(drows, "wf2cu_fold: null pointer");
  Geom g;
  int rc = make_geom(geom, &g, "wf2cu_fold");
  if (rc) return rc;
  View4W dv;
  const float* p;
  rc = view4(dx, "wf2cu_fold", g.N, g.C, g.H, g.W, &p, dv.s);
  if (rc) return rc;
  rc = writable4(dx, "wf2cu_fold");
  if (rc) return rc;
  dv.p = const_cast<float*>(p);
  const int64_t total = (int64_t)g.N * g.C * g.H * g.W;
  if (total == 0) return 0;
  const int64_t slab = (int64_t)g.Ho * g.Wo * g.KH * g.KW;
  if (slab * 4 <= 40 * 1024 && (int64_t)g.N * g.C < (int64_t(1) << 31) && (int64_t)g.H * g.W >= 32) {
    fold_rows_slab_kernel<<<(unsigned)(g.N * g.C), 256, (size_t)slab * 4, S().stream>>>((const float*)drows + drows_off, dv, g, accumulate);
    return post_launch("wf2cu_fold(slab)");
  }
  fold_rows_kernel<<<grid_for(total), 256, 0, S().stream>>>((const float*)drows + drows_off, dv, g, total, accumulate);
  return post_launch("wf2cu_fold");
}

extern "C" int wf2cu_pool2d(int mode, const wf2cu_view* x, const wf2cu_view* out, const wf2cu_unfold_geom* geom) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(mode == WF2CU_POOL_MAX || mode == WF2CU_POOL_AVG, "wf2cu_pool2d: bad mode %d", mode);
  Geom g;
  int rc = make_geom(geom, &g, "wf2cu_pool2d");
  if (rc) return rc;
  View4 xv;
  View4W ov;
  const float* p;
  rc = view4(x, "wf2cu_pool2d(x)", g.N, g.C, g.H, g.W, &xv.p, xv.s);
  if (rc) return rc;
  rc = view4(out, "wf2cu_pool2d(out)", g.N, g.C, g.Ho, g.Wo, &p, ov.s);
  if (rc) return rc;
  rc = writable4(out, "wf2cu_pool2d(out)");
  if (rc) return rc;
  ov.p = const_cast<float*>(p);
  const int64_t total = (int64_t)g.N * g.C * g.Ho * g.Wo;
  if (total == 0) return 0;
  const bool cfast = xv.s[1] == 1 && g.C > 1;
  const unsigned grid = grid_for(total);
  cudaStream_t st = S().stream;
  if (mode == WF2CU_POOL_AVG) {
    if (cfast) pool2d_kernel<true, true><<<grid, 256, 0, st>>>(xv, ov, g, total);
    else pool2d_kernel<true, false><<<grid, 256, 0, st>>>(xv, ov, g, total);
  } else {
    if (cfast) pool2d_kernel<false, true><<<grid, 256, 0, st>>>(xv, ov, g, total);
    else pool2d_kernel<false, false><<<grid, 256, 0, st>>>(xv, ov, g, total);
  }
  return post_launch("wf2cu_pool2d");
}

extern "C" int wf2cu_pool2d_backward(int mode, const wf2cu_view* x, const wf2cu_view* out, const wf2cu_view* dout,
                                     const wf2cu_view* dx, const wf2cu_unfold_geom* geom, int accumulate, int relu_input) {
  WF2_REQUIRE_INIT();
  WF2_CHECK(mode == WF2CU_POOL_MAX || mode == WF2CU_POOL_AVG, "wf2cu_pool2d_backward: bad mode %d", mode);
  Geom g;
  int rc = make_geom(geom, &g, "wf2cu_pool2d_backward");
  if (rc) return rc;
  View4 xv{nullptr, {0, 0, 0, 0}}, ov{nullptr, {0, 0, 0, 0}}, dv;
  View4W gx;
  const float* p;
  if (mode == WF2CU_POOL_MAX || relu_input) {
    rc = view4(x, "wf2cu_pool2d_backward(x)", g.N, g.C, g.H, g.W, &xv.p, xv.s);
    if (rc) return rc;
  }
  if (mode == WF2CU_POOL_MAX) {
    rc = view4(out, "wf2cu_pool2d_backward(out)", g.N, g.C, g.Ho, g.Wo, &ov.p, ov.s);
    if (rc) return rc;
  }
  rc = view4(dout, "wf2cu_pool2d_backward(dout)", g.N, g.C, g.Ho, g.Wo, &dv.p, dv.s);
  if (rc) return rc;
  rc = view4(dx, "wf2cu_pool2d_backward(dx)", g.N, g.C, g.H, g.W, &p, gx.s);
  if (rc) return rc;
  rc = writable4(dx, "wf2cu_pool2d_backward(dx)");
  if (rc) return rc;
  gx.p = const_cast<float*>(p);
  const int64_t total = (int64_t)g.N * g.C * g.H * g.W;
  if (total == 0) return 0;
  const bool cfast = gx.s[1] == 1 && g.C > 1;
  cudaStream_t st = S().stream;
  {
    // windows per pixel <= ceil(KH/sh) * ceil(KW/sw)
    const int64_t wmax = (int64_t)((g.KH + g.sh - 1) / g.sh) * ((g.KW + g.sw - 1) / g.sw);
    const bool need_x = mode == WF2CU_POOL_MAX || relu_input;
    const bool cl = gx.s[1] == 1 && dv.s[1] == 1 && (!need_x || xv.s[1] == 1) && (mode != WF2CU_POOL_MAX || ov.s[1] == 1);
    if (cl && wmax <= 16 && g.C >= 2) {
      const int64_t pixels = (int64_t)g.N * g.H * g.W;
      if (mode == WF2CU_POOL_AVG) pool2d_backward_cl_kernel<true><<<grid_for(pixels), 256, 0, st>>>(xv, ov, dv, gx, g, pixels, accumulate, relu_input);
      else pool2d_backward_cl_kernel<false><<<grid_for(pixels), 256, 0, st>>>(xv, ov, dv, gx, g, pixels, accumulate, relu_input);
      return post_launch("wf2cu_pool2d_backward(channels-last)");
    }
  }
  const unsigned grid = grid_for(total);
  if (mode == WF2CU_POOL_AVG) {
    if (cfast) pool2d_backward_kernel<true, true><<<grid, 256, 0, st>>>(xv, ov, dv, gx, g, total, accumulate, relu_input);
    else pool2d_backward_kernel<true, false><<<grid, 256, 0, st>>>(xv, ov, dv, gx, g, total, accumulate, relu_input);
  } else {
    if (cfast) pool2d_backward_kernel<false, true><<<grid, 256, 0, st>>>(xv, ov, dv, gx, g, total, accumulate, relu_input);
    else pool2d_backward_kernel<false, false><<<grid, 256, 0, st>>>(xv, ov, dv, gx, g, total, accumulate, relu_input);
  }
  return post_launch("wf2cu_pool2d_backward");
}
