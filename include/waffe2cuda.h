/*
 * waffe2cuda.h — C ABI of libwaffe2cuda.so, the B200 (sm_100a) device library behind
 * cl-waffe2's `CUDATensor` backend.
 *
 * This is the drop-in boundary: what `source/backends/cuda/*.lisp` binds through CFFI
 * (`cl-waffe2_b200/lisp/`), exactly where the reference binds `cl-waffe2-simd.so`
 * (`cl-waffe2-simd/api.lisp:63-274`) and Fortran-ABI BLAS
 * (`source/backends/cpu/blas-functions.lisp:8-38`).
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, no C++/torch types.
 *  - every entry point returns int: 0 = ok, non-zero = error; the message is in
 *    wf2cu_last_error() (thread-local).  The reference's C kernels return void and have no
 *    error channel (`cl-waffe2-simd.c:153-155`); the Lisp side turns a non-zero status into a
 *    condition so the VM's `runtime-error` wrapper (`source/vm/vm.lisp:230-256`) shows it.
 *  - all work is enqueued asynchronously on ONE stream per process; the only host syncs are
 *    wf2cu_d2h, wf2cu_get_f32, wf2cu_sync and the timer read-out.
 *  - compute dtype is fp32 (the reference's `:float`); `:uint32` appears only as ArgMax output.
 *  - there is no CPU fallback: without a B200 wf2cu_init fails and every entry point after it
 *    returns WF2CU_ERR_NOT_INITIALISED.
 */
#ifndef WAFFE2CUDA_H
#define WAFFE2CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WF2CU_MAX_RANK 8

/* status codes */
enum {
  WF2CU_OK = 0,
  WF2CU_ERR_CUDA = 1,            /* a CUDA runtime call failed                      */
  WF2CU_ERR_INVALID = 2,         /* bad argument (rank, shape mismatch, null, ...)   */
  WF2CU_ERR_UNSUPPORTED = 3,     /* well-formed request this library does not cover  */
  WF2CU_ERR_NOT_INITIALISED = 4, /* wf2cu_init has not succeeded in this process     */
  WF2CU_ERR_COMM = 5             /* NCCL failure / communicator missing              */
};

/*
 * Strided view descriptor.  Lowered from an AbstractTensor
 * (`source/vm/generic-tensor/tensor.lisp:40-111`): element (i0..i_{r-1}) of the view lives at
 *     base[ offset + sum_k stride[k] * i_k ]          (element units, fp32)
 * where the host has already folded the reference's view algebra into the numbers:
 *     stride[k] = tensor-stride_k * (range_nth(view_k,1) - range_nth(view_k,0))   (0 on :broadcast)
 *     offset    = tensor-initial-offset + sum_k tensor-stride_k * range_nth(view_k,0)
 * (`do-compiled-loop.lisp:170-298`, `iterator/range.lisp:72-96`).  Strides may be 0
 * (broadcast) or negative (reversed slice).  This replaces the per-row
 * `ViewInstruction{offset,size,by}` a reference leaf receives (`view.lisp:50-69`).
 */
typedef struct wf2cu_view {
  void*   base;
  int64_t offset;
  int32_t rank;
  int32_t reserved;
  int64_t shape[WF2CU_MAX_RANK];
  int64_t stride[WF2CU_MAX_RANK];
} wf2cu_view;

/* ---- binary arithmetic: replaces waffe2_s{add,sub,mul,div} / scopy_
 *      (`cl-waffe2-simd.c:153-250`, `backends/cpu/arithmetic.lisp:65-142`) ---- */
enum { WF2CU_ADD = 0, WF2CU_SUB = 1, WF2CU_MUL = 2, WF2CU_DIV = 3, WF2CU_COPY = 4 };

/* ---- unary math: replaces waffe2_s{sin..sqrt} (`cl-waffe2-simd.c:586-706`),
 *      nodes of `base-impl/mathematics.lisp:13-257` ---- */
enum {
  WF2CU_SIN = 0, WF2CU_COS, WF2CU_TAN, WF2CU_ASIN, WF2CU_ACOS, WF2CU_ATAN,
  WF2CU_SINH, WF2CU_COSH, WF2CU_TANH, WF2CU_ABS, WF2CU_EXP, WF2CU_LOGE,
  WF2CU_LOG2, WF2CU_LOG10, WF2CU_LOG1P, WF2CU_SQRT, WF2CU_CBRT,
  WF2CU_SIGN, WF2CU_SQUARE,           /* LispTensor-only nodes in the reference */
  WF2CU_NEG, WF2CU_RECIP,             /* helpers used by fused programs          */
  WF2CU_UNARY_COUNT
};

/* ---- compare: replaces waffe2_s{eq,lt,le,gt,ge}[ _scal ] (`cl-waffe2-simd.c:385-576`) ---- */
enum { WF2CU_EQ = 0, WF2CU_LT = 1, WF2CU_LE = 2, WF2CU_GT = 3, WF2CU_GE = 4 };

/* ===================== runtime ===================== */
int  wf2cu_init(int device);                 /* select device, create the stream; fails without sm_100 */
int  wf2cu_shutdown(void);
const char* wf2cu_last_error(void);
int  wf2cu_device_count(int* out);
int  wf2cu_device_name(char* buf, size_t buflen);
int  wf2cu_sync(void);
int  wf2cu_stream(void** out_cudaStream);    /* the library's stream, for event timing by callers */
int  wf2cu_set_stream(void* cudaStream);     /* adopt a caller-owned stream (NULL = back to own) */
int  wf2cu_launch_count(int64_t* out);       /* kernels launched by this library since init      */

/* storage: the `vec` of a CUDATensor.  Zero-filled like LispTensor's
 * `:initial-element 0` (`backends/lisp/tensor.lisp:20-28`); refcounted because the VM aliases
 * storage with (setf (tensor-vec a) (tensor-vec b)) (`vm/optimize-ir.lisp:4-18`). */
int  wf2cu_malloc(size_t bytes, void** out);
int  wf2cu_retain(void* dev);
int  wf2cu_release(void* dev);               /* frees at refcount 0 */
int  wf2cu_free(void* dev);                  /* == release           */
int  wf2cu_memset0(void* dev, size_t offset_bytes, size_t bytes);
int  wf2cu_h2d(void* dev, size_t offset_bytes, const void* host, size_t bytes);
int  wf2cu_d2h(const void* dev, size_t offset_bytes, void* host, size_t bytes);   /* syncs */
/* enqueue only, on the current (branch) stream: `host` must be pinned (wf2cu_host_alloc) and may be read after wf2cu_sync,
 * wf2cu_timer_stop or a wf2cu_d2h on a stream that has joined the copy's branch.  Bulk results of a step leave the device this
 * way while the next step's kernels run (the reference's results are already in host memory, `array-converter.lisp:154-218`). */
int  wf2cu_d2h_async(const void* dev, size_t offset_bytes, void* host, size_t bytes);
int  wf2cu_d2d(void* dst, size_t dst_off, const void* src, size_t src_off, size_t bytes);
int  wf2cu_get_f32(const void* dev, int64_t index, float* out);                    /* vref, syncs */
int  wf2cu_set_f32(void* dev, int64_t index, float value);                         /* (setf vref) */
int  wf2cu_host_alloc(size_t bytes, void** out);   /* pinned staging for h2d/d2h */
int  wf2cu_host_free(void* host);
int  wf2cu_mem_in_use(size_t* bytes);

/* timing on the library stream (CUDA events) */
int  wf2cu_timer_start(void);
int  wf2cu_timer_stop(float* elapsed_ms);    /* syncs on the stop event */
int  wf2cu_l2_flush(void);                   /* writes a >L2 scratch buffer */

/* CUDA-graph capture of a static instruction sequence (`vm/vm.lisp:260-289`) */
int  wf2cu_graph_begin(void);
int  wf2cu_graph_end(void** out_exec);
int  wf2cu_graph_launch(void* exec);
int  wf2cu_graph_destroy(void* exec);
/* Side branches of a compiled step (replaces nothing in the reference: its VM runs the instruction
 * sequence in order, `vm/vm.lisp:260-289`; independent instructions overlap here).  begin(b), b = 1..3:
 * following launches go to side stream b, ordered after everything issued so far; end(b): back to the main
 * stream; join(b): the main stream waits for branch b.  Captured as fork/join edges by wf2cu_graph_begin/end. */
int  wf2cu_branch_begin(int b);
int  wf2cu_branch_end(int b);
int  wf2cu_branch_join(int b);

/* ===================== element-wise ===================== */
/* dst <- dst OP src over views of equal shape.  Axes where dst has stride 0 and extent > 1 are
 * REDUCTION axes: the reference runs `x[0] = x[0] OP y[i]` sequentially there
 * (`cl-waffe2-simd.c:159,205-212`); here they are classified and tree-reduced
 * (ADD, SUB, MUL; COPY keeps the last element).  This is how `!sum` arrives
 * (`base-impl/reduction.lisp:60-79`). */
int  wf2cu_ew_binary(int op, const wf2cu_view* dst, const wf2cu_view* src);
/* out <- a OP b, the fused form of MoveTensorNode + {Add,Sub,Mul,Div}Node
 * (`base-impl/arithmetic.lisp:150-152`).  out must not have reduction axes. */
int  wf2cu_ew_binary3(int op, const wf2cu_view* out, const wf2cu_view* a, const wf2cu_view* b);
/* dst <- dst OP scalar: the `A*=scal` family without the 4-byte H2D of `Scalar->MatNode`
 * (`base-impl/fundamental.lisp:525-533`). */
int  wf2cu_ew_scalar(int op, const wf2cu_view* dst, float scalar);
int  wf2cu_fill(const wf2cu_view* dst, float value);
/* dst <- f(src) */
int  wf2cu_ew_unary(int fn, const wf2cu_view* src, const wf2cu_view* dst);
/* dst <- src ^ n  (ExptNode, `backends/cpu/mathematics.lisp:55-75`) */
int  wf2cu_pow(const wf2cu_view* src, const wf2cu_view* dst, float n);
/* dst <- (x CMP y) ? then : else ; y == NULL compares against `scal`
 * (`backends/cpu/logical.lisp:37-84`) */
int  wf2cu_cmp(int op, const wf2cu_view* x, const wf2cu_view* y, float scal,
               const wf2cu_view* dst, float then_value, float else_value);

/* Fused element-wise chain: one pass over `nviews` equal-shape views running a register
 * program.  Each instruction is 4 x int32: {opcode, dst_reg, a_reg, b_reg_or_imm_bits}.
 * See WF2CU_P_* below.  Up to 16 registers, 64 instructions, 8 views. */
enum {
  WF2CU_P_LOAD = 0,   /* r[dst] = view[a]                          */
  WF2CU_P_STORE,      /* view[a] = r[dst]                          */
  WF2CU_P_IMM,        /* r[dst] = bits(b) as float                 */
  WF2CU_P_ADD, WF2CU_P_SUB, WF2CU_P_MUL, WF2CU_P_DIV,   /* r[dst] = r[a] op r[b] (rn, no fma contraction) */
  WF2CU_P_UNARY,      /* r[dst] = f_b(r[a]), b = WF2CU_SIN...      */
  WF2CU_P_POW,        /* r[dst] = pow(r[a], r[b])                  */
  WF2CU_P_CMPSEL,     /* r[dst] = (r[a] CMP r[b]) ? 1 : 0 ; CMP in bits 8.. of opcode */
  WF2CU_P_MOV         /* r[dst] = r[a]                             */
};
int  wf2cu_fused(const int32_t* program, int n_instr, const wf2cu_view* views, int nviews);
/* Executors of wf2cu_fused.  The programs the host fusion passes emit (sigmoid, relu and its backward, gelu, the MSE and
 * softmax-cross-entropy chains, SGD on views, sin(sin(sin x)) of `benchmarks/compose.lisp:21-26`) are compiled kernels
 * (path 2: straight-line code, HBM-bound); anything else, and strided or misaligned views, runs on the per-element
 * interpreter (path 1).  Both apply the same operations in the same order: results are bit-identical.
 * set_mode(1) forces the interpreter (used by the parity tests); last_path: 0 none yet, 1 interpreter, 2 compiled. */
int  wf2cu_fused_set_mode(int mode);
int  wf2cu_fused_last_path(int* out);
int  wf2cu_fused_catalogue_size(int* out);

/* ===================== reductions / composites ===================== */
/* out <- [out +] sum(x) over the axes where out has stride 0 or extent 1 vs x; the fused form of
 * `A*=scal out 0` + `A+=B out* x` (`base-impl/reduction.lisp:60-79`).  scale multiplies the
 * result (1/count gives `!mean`, `reduction.lisp:81-136`; division is done as a true divide by
 * `divisor` when divisor != 0 to match `!div`). */
int  wf2cu_reduce_sum(const wf2cu_view* x, const wf2cu_view* out_bcast, int accumulate, float divisor);
/* Allocation-free `sum :axis t`.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * out[out_off] <- [out[out_off] +] sum(x[x_off .. x_off + n)) [/ divisor]  (fp32 element offsets, device pointers,
 * x + x_off 16-byte aligned), on the library's stream — the flat-sum kernel of wf2cu_reduce_sum with its per-CTA
 * partials and last-block ticket in a file-scope `static __device__` array of 4352 bytes (1024 partials + one 256-byte
 * ticket line) that every GPU loading the library holds until exit.  It calls no cudaMalloc / cudaFree / allocating
 * routine.  n > WF2CU_NOALLOC_MAX_ELEMS is refused with WF2CU_ERR_UNSUPPORTED (use wf2cu_reduce_sum).  Every call
 * shares the one array (the ticket is cleared on the stream before each launch, the partials are overwritten before
 * they are read): calls that overlap on the GPU — other streams, other host threads — race; serialise them. */
#define WF2CU_NOALLOC_MAX_ELEMS (1ll << 28)   /* 1 GiB of fp32 */
int  wf2cu_sum_all_noalloc(const void* x, int64_t x_off, int64_t n, void* out, int64_t out_off, int accumulate, float divisor);
/* `!softmax` (`nn/activation.lisp:126-159`): mean-shift (avoid_overflow) or plain, over `axis`. */
int  wf2cu_softmax(const wf2cu_view* x, const wf2cu_view* out, int axis, int avoid_overflow);
/* `(!softmax (!div x n))`: the scalar divide in front folded into the same pass (GPT-2's attention weights,
 * `examples/gpt-2/utils.lisp:236-242`).  pre_divisor = 0: plain wf2cu_softmax.  Implemented for a softmax axis that
 * is not the innermost one; WF2CU_ERR_UNSUPPORTED otherwise (issue the two nodes separately). */
int  wf2cu_softmax_ex(const wf2cu_view* x, const wf2cu_view* out, int axis, int avoid_overflow, float pre_divisor);
/* MaxValue/MinValue/ArgMax/ArgMin over the last axis (`backends/cpu/matrix-ops.lisp:240-369`).
 * out has x's shape with the last axis of extent 1; arg variants write uint32 indices. */
int  wf2cu_reduce_maxmin(int is_min, const wf2cu_view* x, const wf2cu_view* out);
int  wf2cu_argmaxmin(int is_min, const wf2cu_view* x, const wf2cu_view* out_u32);

/* ===================== MatmulNode ===================== */
/* C[b] <- op(A[b]) . op(B[b]), row-major, alpha=1 beta=0, replaces the reference's
 * `sgemm_` call (`backends/cpu/matrix-ops.lisp:58-79`): op(A) is m x k, op(B) is k x n,
 * C is m x n; trans flag = the operand is stored transposed (`LazyTransposeNode`).
 * ld* in elements; batch strides in elements (0 = shared operand).
 * Numerics: fp32-faithful (3xTF32 split on tcgen05, or exact fp32 FMA for shapes the tensor
 * path does not take). */
int  wf2cu_gemm(int trans_a, int trans_b, int64_t m, int64_t n, int64_t k,
                const void* A, int64_t a_off, int64_t lda,
                const void* B, int64_t b_off, int64_t ldb,
                void* C, int64_t c_off, int64_t ldc,
                int64_t batch, int64_t stride_a, int64_t stride_b, int64_t stride_c);
/* 0 = auto, 1 = force fp32 SIMT path, 2 = force tcgen05 3xTF32 (planes + persistent 128x256-tile kernel),
 * 3 = tcgen05 1xTF32 (not parity-grade), 4 = force the one-launch fused tcgen05 kernel (small problems),
 * 5 / 6 = force tcgen05 3xTF32 on the persistent 256x256 / 256x192-tile kernel, 7 = force the raw-operand kernel (A not
 * transposed: fp32 tiles straight from the operand by TMA, lo planes made on the SM, no pre-pass), 8 = force the CTA-pair kernel
 * (tcgen05.mma.cta_group::2 on 512x256 pair tiles: each CTA of a pair loads half of the B tile).  Forced modes return
 * WF2CU_ERR_UNSUPPORTED for shapes their kernel does not take. */
int  wf2cu_gemm_set_mode(int mode);
/* How the plane-fed tcgen05 kernels form the error-compensation terms of the split product
 *   C = A_hi.B_hi + (A_hi.B_lo + A_lo.B_hi),  A_hi = tf32_rn(A), A_lo = A - A_hi:
 * BF16 (default): the two correction products run as kind::f16 MMAs on bf16_rn(A), bf16_rn(A_lo) (K = 16 per instruction,
 * twice the TF32 rate): 2 tensor-pipe units per k instead of 3.  The corrections are ~2^-12 of the product, so rounding their
 * factors to bf16 leaves ~5e-7 of max|C| on random operands — the size of an fp32 sgemm's own rounding error (the reference
 * leaf is OpenBLAS sgemm_, `source/backends/cpu/matrix-ops.lisp:58-79`); small integers stay exact (lo = 0).
 * TF32: the classic 3xTF32 form (corrections as kind::tf32 MMAs on tf32_rn(A_lo)), ~6e-8.  Start-up default from
 * WF2CU_GEMM_CORRECTION=tf32.  The raw-operand and small-problem kernels always use the TF32 form. */
enum { WF2CU_GEMM_CORRECTION_BF16 = 0, WF2CU_GEMM_CORRECTION_TF32 = 1 };
int  wf2cu_gemm_set_correction(int kind);
/* Which kernel the last wf2cu_gemm / wf2cu_gemm_ex call launched (the parity tests assert that the benchmarked shapes
 * really run on the kernel whose throughput is quoted). */
enum {
  WF2CU_GEMM_PATH_NONE = 0,
  WF2CU_GEMM_PATH_SIMT = 1,          /* fp32 FMA tile kernel                                   */
  WF2CU_GEMM_PATH_SIMT_SPLITK = 2,   /* fp32 FMA, K over a cluster                             */
  WF2CU_GEMM_PATH_SIMT_SKINNY = 3,   /* fp32 FMA, Conv2D's tall / long-K shapes                */
  WF2CU_GEMM_PATH_TC_FUSED = 4,      /* tcgen05 3xTF32, one launch, split in registers         */
  WF2CU_GEMM_PATH_TC_128X256 = 5,    /* tcgen05 3xTF32, hi/lo planes + persistent 128x256 tiles */
  WF2CU_GEMM_PATH_TC_256X256 = 6,    /* ... 256x256 tiles                                       */
  WF2CU_GEMM_PATH_TC_256X192 = 7,    /* ... 256x192 tiles                                       */
  WF2CU_GEMM_PATH_TC_RAW = 8,        /* tcgen05 3xTF32, raw fp32 tiles by TMA, lo planes made on the SM (no pre-pass) */
  WF2CU_GEMM_PATH_TC_2SM = 9         /* hi/lo planes + persistent 512x256 pair tiles on CTA pairs (cta_group::2)       */
};
int  wf2cu_gemm_last_path(int* out);
/* Prepared operands — the packed-operand idea of BLAS libraries (sgemm_pack / sgemm_compute) for the leaf the reference calls
 * (`sgemm_`, `source/backends/cpu/matrix-ops.lisp:58-79`, which re-reads its operands every call).  The host declares
 * count fp32 elements from base + offset constant (constant != 0): the parameters of a model that is only run forward
 * (`examples/gpt-2/gpt2-model.lisp`).  A MatmulNode operand that STARTS inside such a range keeps its tensor-core operand
 * planes between calls (found again by pointer, pitch, shape and layout), so its pre-pass runs once.  wf2cu_h2d / wf2cu_d2d /
 * wf2cu_memset0 into the range mark the planes stale (split again by the next product); wf2cu_free of the storage, or
 * constant == 0, drops them.  Kernels of this library that write into a declared range are the caller's contract
 * violation.  Planes are allocated on first use: not inside wf2cu_graph_begin/end (run the product once first).
 * stats: number of kept plane sets and their bytes. */
int  wf2cu_gemm_const_operand(const void* base, int64_t offset, int64_t count, int constant);
int  wf2cu_gemm_const_stats(int64_t* entries, int64_t* bytes);
/* The result of one node is very often the K-major A operand of the MatmulNode that follows it (LinearLayer after LayerNorm,
 * `nn/regression.lisp:14-18` after `examples/gpt-2/utils.lisp:114-127`; the projections of `examples/gpt-2/model.lisp:93-137`).
 * The product's operand pre-pass then only restates what the producing kernel had in registers.  on != 0: the NEXT call of
 * wf2cu_layernorm, wf2cu_attn_pv / wf2cu_attn_fused, or wf2cu_gemm / _ex / _ex2 (one shot: a wf2cu_gemm* call consumes the request
 * whether or not it can honour it — it can on the 256-row tensor-core kernels, dense unbatched C with n a multiple of 32; other
 * entry points leave it pending) also writes the tensor-core operand plane of its result; a
 * wf2cu_gemm / _ex / _ex2 call that comes IMMEDIATELY after it (no other launch of this library in between) and reads exactly
 * that result as its untransposed A operand uses the plane instead of running the pre-pass.  Same bits as the pre-pass writes,
 * so the product is bit-identical either way; a producer or a shape that cannot (k not a multiple of 32, TF32 corrections
 * selected, inside a graph capture before the buffer exists, a side branch) ignores the request. */
int  wf2cu_emit_operand_planes(int on);
/* MatmulNode with the node that follows it folded into the epilogue (what the backend's
 * `on-finalizing-compiling` pass emits for LinearLayer, `nn/regression.lisp:14-18`, and for the
 * `!relu` that follows / precedes it, `nn/activation.lisp:11-26`):
 *   C = op(A) op(B) [+ bias[col]]  ;  act = 1: C = C * (C > 0 ? 1 : 0)  ;  mask != NULL: C = C * (mask[row,col] > 0 ? 1 : 0)
 * (the last form is the MulNode backward of `!relu` applied to dX = dout . W).  Unbatched. */
int  wf2cu_gemm_ex(int trans_a, int trans_b, int64_t m, int64_t n, int64_t k,
                   const void* A, int64_t a_off, int64_t lda, const void* B, int64_t b_off, int64_t ldb,
                   void* C, int64_t c_off, int64_t ldc, const void* bias, int64_t bias_off, int act,
                   const void* mask, int64_t mask_off, int64_t ldmask);
/* The GPT-2 example's block (`examples/gpt-2/model.lisp:93-137`) follows two of its projections with one more node: `!gelu`
 * after the first MLP projection (`utils.lisp:129-137`) and the residual `(!add x ...)` after the attention and MLP output
 * projections.  Both ride in the epilogue here:  C = op(A) op(B) [+ bias[col]] ; act = 1: relu, act = 2: gelu (tanh form, the
 * operations and order of the node) ; residual != NULL: C = residual[row, col] + C.  Bit-identical to the separate nodes.
 * Unbatched. */
int  wf2cu_gemm_ex2(int trans_a, int trans_b, int64_t m, int64_t n, int64_t k,
                    const void* A, int64_t a_off, int64_t lda, const void* B, int64_t b_off, int64_t ldb,
                    void* C, int64_t c_off, int64_t ldc, const void* bias, int64_t bias_off, int act,
                    const void* residual, int64_t res_off, int64_t ldres);
/* `(!sum (softmax-cross-entropy x labels))` (`nn/criterion.lisp:89-117,164-188`; the MNIST example's
 * loss, `examples/mnist/mlp.lisp:24-28`) in one kernel: z = softmax(x) row-wise (avoid-overflow nil),
 * loss[0] = sum(-1 * labels * log(z + delta)); x, labels, z contiguous (rows cols), cols <= 1024.
 * wf2cu_sce_backward is `sce-diff-static` (:81-87): dx = (dy[0] * (z - labels)) / rows. */
int  wf2cu_sce_sum_forward(const void* x, int64_t x_off, const void* labels, int64_t l_off, void* z, int64_t z_off,
                           void* loss, int64_t loss_off, int64_t rows, int64_t cols, float delta);
int  wf2cu_sce_backward(const void* z, int64_t z_off, const void* labels, int64_t l_off, const void* dy, int64_t dy_off,
                        void* dx, int64_t dx_off, int64_t rows, int64_t cols, int accumulate);
/* Forward and `sce-diff-static` in ONE launch, for an executor that already knows the loss' incoming gradient when it issues the
 * forward (a training step seeds the summed loss with the constant 1): dx = (dy[0] * (z - labels)) / rows is written beside z, with
 * the operations of wf2cu_sce_backward (accumulate = 0), so both are bit-identical to the two calls.  dy = dx = NULL: the plain forward. */
int  wf2cu_sce_sum_forward_grad(const void* x, int64_t x_off, const void* labels, int64_t l_off, void* z, int64_t z_off,
                                void* loss, int64_t loss_off, int64_t rows, int64_t cols, float delta,
                                const void* dy, int64_t dy_off, void* dx, int64_t dx_off);

/* ===================== optimizers (fused chains of `optimizers/impls/*.lisp`) ===================== */
/* p <- p - lr * g        (`sgd.lisp:20-26`) */
int  wf2cu_sgd_step(void* param, int64_t p_off, const void* grad, int64_t g_off, int64_t n, float lr);
/* Adam as written in `adam.lisp:45-116`:  m <- m + (g - m)(1 - b1) ; v <- v + (1 - b2)(g^2 - v) ;
 * p <- p - (m * lr_t) / (eps + sqrt(v)), lr_t computed by the host (`adam-step-lr`).  When
 * lr_t_dev != NULL the kernel reads lr_t from that device float instead (so a captured CUDA graph
 * can be replayed with a fresh value).  grad_scale pre-multiplies g (1/world for the data-parallel
 * mean). */
int  wf2cu_adam_step(void* param, int64_t p_off, const void* grad, int64_t g_off,
                     void* m, int64_t m_off, void* v, int64_t v_off, int64_t n,
                     float lr_t, const void* lr_t_dev, float beta1, float beta2, float eps, float grad_scale);

/* `LayerNorm-Revisit` (`examples/gpt-2/utils.lisp:114-127`) over the last axis of dense rows, the ten-node chain
 * (!mean, !sub, !expt 2, !mean, !add eps, !sqrt, !div, !mul g, !add b) as one pass with the same per-element
 * operations in the same order.  cols must be a multiple of 4, <= 1024, buffers 16-byte aligned; otherwise
 * WF2CU_ERR_UNSUPPORTED (issue the nodes). */
int  wf2cu_layernorm(const void* x, int64_t x_off, const void* g, int64_t g_off, const void* b, int64_t b_off,
                     void* out, int64_t out_off, int64_t rows, int64_t cols, float eps);

/* `GPT2PositionalEmbedding` (`examples/gpt-2/utils.lisp:19-112`):
 * out[b,p,:] = wte[round(ctx[b,p]),:] + wpe[p,:] ; token ids are floats. */
int  wf2cu_embedding_gather(const void* ctx, int64_t batch, int64_t seq, const void* wte,
                            const void* wpe, int64_t dim, void* out);

/* ===================== the GPT-2 example's SelfAttention ===================== */
/* `SelfAttention` of `examples/gpt-2/utils.lisp:204-244`, as written there (scale 1 / N with N = the last axis of the score
 * tensor, `!softmax :avoid-overflow nil` over its default axis 1 = the HEADS axis of the 4-D tensor, no causal mask), in two
 * launches instead of split-heads copies + `!matmul` + `!div` + `!softmax` + `!matmul` + merge-heads copies:
 *   wf2cu_attn_heads_softmax : p[b,h,i,j] = e / sum_h' e[b,h',i,j],  e = exp((Q_h[i,:] . K_h[j,:]) / divisor)
 *   wf2cu_attn_pv            : out[b,i,h*dim:(h+1)*dim] = sum_j p[b,h,i,j] V_h[j,:]        (merged heads, as merge-heads leaves them)
 * qkv: the dense (batch seq 3*n_head*head_dim) activation of the fused QKV projection (Q | K | V along the last axis, heads
 * consecutive inside each); p: dense (batch n_head seq seq); out: dense (batch seq n_head*head_dim).
 * Numerics: both products are the error-compensated 3xTF32 of MatmulNode (raw fp32 tiles are the hi planes, lo = x - trunc(x)).
 * Implemented for n_head = 12, head_dim = 64, seq a multiple of 128, 16-byte aligned operands; WF2CU_ERR_UNSUPPORTED otherwise
 * (issue the nodes separately, exactly as the example does). */
int  wf2cu_attn_heads_softmax(const void* qkv, int64_t qkv_off, int64_t batch, int64_t seq, int64_t n_head, int64_t head_dim,
                              float divisor, void* p, int64_t p_off);
int  wf2cu_attn_pv(const void* p, int64_t p_off, const void* qkv, int64_t qkv_off, int64_t batch, int64_t seq, int64_t n_head,
                   int64_t head_dim, void* out, int64_t out_off);

/* The same block as ONE call when the attention weights themselves are not wanted (the example's `SelfAttention` returns only the
 * merged-heads result): the score kernel makes ONE pass over the 12 heads, writes the UN-normalised e = exp(s / divisor) and the
 * reciprocal denominators r[b,i,j] = RN(1 / sum_h e) into a workspace of the library, and the P.V kernel forms p = RN(e r) — the very
 * product wf2cu_attn_heads_softmax forms — on the way into tensor memory.  Half the Q.K^T work and one exp per element instead of
 * two; out is bit-identical to wf2cu_attn_heads_softmax + wf2cu_attn_pv.  Workspace: (batch n_head seq seq) + (batch seq seq) floats,
 * grow-only, allocated on first use (not inside wf2cu_graph_begin/end: run once first).  Same shape limits and status codes. */
int  wf2cu_attn_fused(const void* qkv, int64_t qkv_off, int64_t batch, int64_t seq, int64_t n_head, int64_t head_dim, float divisor,
                      void* out, int64_t out_off);

/* ===================== unfold / Conv2D / pooling ===================== */
/* Geometry of one unfold, the argument list of `waffe2_im2col_s_column` / `waffe2_col2im_s_column`
 * (`cl-waffe2-simd/kernels/im2col.c:63-80`; CFFI `cl-waffe2-simd/api.lisp:27-61`; slots of
 * `Im2ColNode`, `source/base-impl/unfold.lisp:9-23`).  For wf2cu_im2col / wf2cu_col2im H, W are the
 * extents of the already padded image the node receives; for the fused entry points below they are
 * the extents of the unpadded input. */
typedef struct wf2cu_unfold_geom {
  int64_t N, C, H, W;
  int64_t Ho, Wo, KH, KW;
  int64_t pad_h, pad_w, stride_h, stride_w, dilation_h, dilation_w;
} wf2cu_unfold_geom;

/* Im2ColNode forward (`source/backends/cpu/im2col.lisp:4-25`): col (N C KH KW Ho Wo) <- im (N C H W),
 * both dense C-order buffers (`:order :column`; the reference's `_row` variants index `col` out of
 * bounds and are not offered).  As in the reference, output rows j with j*stride_h + pad_h >=
 * stride_h*Ho (and columns likewise) are left untouched.  dilation != 1 -> WF2CU_ERR_UNSUPPORTED. */
int  wf2cu_im2col(void* col, int64_t col_off, const void* im, int64_t im_off, const wf2cu_unfold_geom* g);
/* Col2ImNode forward (`source/backends/cpu/im2col.lisp:27-50`): im <- 0 ; im += col over the same walk;
 * per image element the contributions are added in (kh, kw) ascending order, as the C loop does. */
int  wf2cu_col2im(const void* col, int64_t col_off, void* im, int64_t im_off, const wf2cu_unfold_geom* g);

/* `unfold` (`source/nn/im2col.lisp:56-98`) as one pass: zero padding ((pad, pad + Ho mod stride) per
 * axis), Im2ColNode, `(!permute (torch-order 0 4 5 1 2 3))`, `->contiguous`, reshape.
 * x: rank-4 view (N C H W) of the UNPADDED input, any strides; rows: dense (N*Ho*Wo, C*KH*KW).
 * Positions the node leaves untouched are written as 0. */
int  wf2cu_unfold(const wf2cu_view* x, void* rows, int64_t rows_off, const wf2cu_unfold_geom* g);
/* Backward of that chain as one pass (reshape, Permute-Node backward, Col2ImNode, the window of the
 * padded gradient, and the gradient adder `vm/utils.lisp:109-142`): dx (N C H W view) <- or += fold(drows). */
int  wf2cu_fold(const void* drows, int64_t drows_off, const wf2cu_view* dx, const wf2cu_unfold_geom* g, int accumulate);

/* MaxPool2D / AvgPool2D (`source/nn/pool.lisp:73-105`): unfold -> windows (.., KH*KW) -> `!max` / `!mean`
 * over axis 1 -> (N Ho Wo C) -> permute, as one pass.  x: (N C H W) view, out: (N C Ho Wo) view (the
 * reference's result is a permuted view over (N Ho Wo C) storage; any strides are accepted).
 * avg: sequential fp32 sum in window order, then / (KH*KW), as `!mean` does. */
enum wf2cu_pool_mode { WF2CU_POOL_MAX = 0, WF2CU_POOL_AVG = 1 };
int  wf2cu_pool2d(int mode, const wf2cu_view* x, const wf2cu_view* out, const wf2cu_unfold_geom* g);
/* Backward: max = MaxValue-Node backward (`base-impl/matrix-ops.lisp:311-334`: (a == max a) * dout, every
 * tie gets the gradient) followed by the unfold backward; avg = dout / (KH*KW) spread over the window.
 * x and out are read for max only (may be NULL for avg).  dx <- or += .
 * relu_input != 0: x is the result of a fused `... + !relu` node (wf2cu_gemm_ex act 1); the `!relu` backward
 * `dout * (x > 0 ? 1 : 0)` (`nn/activation.lisp:11-26`) is applied to the contribution here (x is then read
 * for avg too). */
int  wf2cu_pool2d_backward(int mode, const wf2cu_view* x, const wf2cu_view* out, const wf2cu_view* dout,
                           const wf2cu_view* dx, const wf2cu_unfold_geom* g, int accumulate, int relu_input);

/* ===================== multi-GPU (batch-sharded MLP step) ===================== */
int  wf2cu_comm_unique_id(void* out128);                       /* rank 0: 128-byte NCCL id   */
int  wf2cu_comm_init(int rank, int world, const void* id128);  /* one process per GPU        */
int  wf2cu_allreduce_sum_f32(void* buf, int64_t offset, int64_t count);  /* on the stream    */
int  wf2cu_comm_destroy(void);

/* Exchange + optimizer as ONE kernel over NVLink peer memory (new work; the reference has no
 * distributed code — the step it replaces is `call-optimizer!` after `backward`,
 * `vm/generic-tensor/acceptor.lisp:466-508`, run once per replica).  One process per GPU:
 *   1. every rank: wf2cu_p2p_export(gradient arena), wf2cu_p2p_export(parameter arena) and
 *      wf2cu_p2p_signal_export() -> three 64-byte CUDA-IPC handles; the host all-gathers them (any transport);
 *   2. wf2cu_p2p_attach(rank, world, local arenas, 3 x (world x 64 B))   (world <= 8);
 *   3. per step, after backward: wf2cu_p2p_allreduce_adam_step — waits until every rank's gradients are
 *      complete; rank r sums shard r of the bucket over all ranks' arenas in rank order, scales by
 *      1/world, applies Adam (same formula as wf2cu_adam_step) with its m/v shard and stores the new
 *      parameters into every replica; then waits until every rank has written its shard, so replicas
 *      stay bit-identical and the gradient arena can be reused by the next backward.
 * Buckets of at most 2^20 floats take the LL (push) form of the same step: every 16-byte store over NVLink carries
 * {value, epoch, value, epoch}, receivers poll their own staging area (part of the window wf2cu_p2p_signal_export exports), so
 * there are no fences, flags or "done" exchange on the path; same summation order and arithmetic, bit-identical results
 * (WF2CU_P2P_LL=0 forces the pull form above; the bucket size must not change within one attachment).
 * A peer that never arrives makes the kernel give up after 10 s (the GPU is never left spinning).  Giving up is sticky and
 * fatal for the attachment: the rank raises its flag (wf2cu_p2p_status: a mapped pinned host word, read without a sync),
 * tells every peer, and from then on every rank's step kernel returns at entry without touching gradients or parameters.
 * wf2cu_p2p_detach resets flags and epochs: the host must pass a barrier across ranks before calling it (every rank's
 * kernels finished), and need not synchronise before the first step of a new attachment.
 * g_off, p_off, m_off, v_off: element offsets, multiples of 4. */
int  wf2cu_p2p_export(void* dev_alloc, void* handle64);
int  wf2cu_p2p_signal_export(void* handle64);
int  wf2cu_p2p_attach(int rank, int world, void* local_grad_alloc, void* local_param_alloc,
                      const void* grad_handles, const void* param_handles, const void* signal_handles);
int  wf2cu_p2p_allreduce_adam_step(void* param, int64_t p_off, int64_t g_off, void* m, int64_t m_off,
                                   void* v, int64_t v_off, int64_t n, float lr_t, const void* lr_t_dev,
                                   float beta1, float beta2, float eps);
int  wf2cu_p2p_status(int* timed_out);
int  wf2cu_p2p_detach(void);

#ifdef __cplusplus
}
#endif
#endif /* WAFFE2CUDA_H */
