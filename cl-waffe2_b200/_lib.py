"""ctypes binding of libwaffe2cuda.so — the Python stand-in for the CFFI `defcfun`s of
`cl-waffe2_b200/lisp/cffi.lisp` (the reference binds its native kernels the same way in
`/root/reference/cl-waffe2-simd/api.lisp:63-274`).  Prototypes mirror include/waffe2cuda.h 1:1.

There is NO fallback: if the shared library is missing this module raises, and if no B200 is
visible `init()` raises.  (The reference instead warns and disables the backend,
`cl-waffe2-simd/shared-object.lisp:3-12`; a device backend without its device has nothing to
fall back to.)"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libwaffe2cuda.so")
MAX_RANK = 8


class Wf2CudaError(RuntimeError):
    """A non-zero status from libwaffe2cuda.so (message from wf2cu_last_error)."""

    def __init__(self, code: int, msg: str):
        super().__init__(f"[wf2cu status {code}] {msg}")
        self.code = code


class View(C.Structure):
    """struct wf2cu_view (include/waffe2cuda.h)."""
    _fields_ = [("base", C.c_void_p), ("offset", C.c_int64), ("rank", C.c_int32), ("reserved", C.c_int32),
                ("shape", C.c_int64 * MAX_RANK), ("stride", C.c_int64 * MAX_RANK)]


class UnfoldGeom(C.Structure):
    """struct wf2cu_unfold_geom (include/waffe2cuda.h)."""
    _fields_ = [(n, C.c_int64) for n in ("N", "C", "H", "W", "Ho", "Wo", "KH", "KW", "pad_h", "pad_w",
                                         "stride_h", "stride_w", "dilation_h", "dilation_w")]


VP = C.POINTER(View)
GP = C.POINTER(UnfoldGeom)
i32, i64, f32, vp, sz = C.c_int, C.c_int64, C.c_float, C.c_void_p, C.c_size_t

# name -> argtypes ; every function returns int status except wf2cu_last_error
PROTOTYPES = {
    "wf2cu_init": [i32], "wf2cu_shutdown": [], "wf2cu_device_count": [C.POINTER(i32)],
    "wf2cu_device_name": [C.c_char_p, sz], "wf2cu_sync": [], "wf2cu_stream": [C.POINTER(vp)],
    "wf2cu_set_stream": [vp], "wf2cu_launch_count": [C.POINTER(i64)],
    "wf2cu_malloc": [sz, C.POINTER(vp)], "wf2cu_retain": [vp], "wf2cu_release": [vp], "wf2cu_free": [vp],
    "wf2cu_memset0": [vp, sz, sz], "wf2cu_h2d": [vp, sz, vp, sz], "wf2cu_d2h": [vp, sz, vp, sz],
    "wf2cu_d2h_async": [vp, sz, vp, sz], "wf2cu_d2d": [vp, sz, vp, sz, sz], "wf2cu_get_f32": [vp, i64, C.POINTER(f32)], "wf2cu_set_f32": [vp, i64, f32],
    "wf2cu_host_alloc": [sz, C.POINTER(vp)], "wf2cu_host_free": [vp], "wf2cu_mem_in_use": [C.POINTER(sz)],
    "wf2cu_timer_start": [], "wf2cu_timer_stop": [C.POINTER(f32)], "wf2cu_l2_flush": [],
    "wf2cu_graph_begin": [], "wf2cu_graph_end": [C.POINTER(vp)], "wf2cu_graph_launch": [vp],
    "wf2cu_graph_destroy": [vp],
    "wf2cu_branch_begin": [i32], "wf2cu_branch_end": [i32], "wf2cu_branch_join": [i32],
    "wf2cu_ew_binary": [i32, VP, VP], "wf2cu_ew_binary3": [i32, VP, VP, VP], "wf2cu_ew_scalar": [i32, VP, f32],
    "wf2cu_fill": [VP, f32], "wf2cu_ew_unary": [i32, VP, VP], "wf2cu_pow": [VP, VP, f32],
    "wf2cu_cmp": [i32, VP, VP, f32, VP, f32, f32], "wf2cu_fused": [C.POINTER(C.c_int32), i32, VP, i32],
    "wf2cu_fused_set_mode": [i32], "wf2cu_fused_last_path": [C.POINTER(i32)], "wf2cu_fused_catalogue_size": [C.POINTER(i32)],
    "wf2cu_reduce_sum": [VP, VP, i32, f32], "wf2cu_sum_all_noalloc": [vp, i64, i64, vp, i64, i32, f32], "wf2cu_softmax": [VP, VP, i32, i32], "wf2cu_softmax_ex": [VP, VP, i32, i32, f32],
    "wf2cu_reduce_maxmin": [i32, VP, VP], "wf2cu_argmaxmin": [i32, VP, VP],
    "wf2cu_gemm": [i32, i32, i64, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, i64, i64, i64, i64],
    "wf2cu_gemm_set_mode": [i32], "wf2cu_gemm_set_correction": [i32], "wf2cu_emit_operand_planes": [i32], "wf2cu_gemm_last_path": [C.POINTER(i32)],
    "wf2cu_gemm_const_operand": [vp, i64, i64, i32], "wf2cu_gemm_const_stats": [C.POINTER(i64), C.POINTER(i64)],
    "wf2cu_gemm_ex": [i32, i32, i64, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i32, vp, i64, i64],
    "wf2cu_gemm_ex2": [i32, i32, i64, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i32, vp, i64, i64],
    "wf2cu_sce_sum_forward": [vp, i64, vp, i64, vp, i64, vp, i64, i64, i64, f32],
    "wf2cu_sce_sum_forward_grad": [vp, i64, vp, i64, vp, i64, vp, i64, i64, i64, f32, vp, i64, vp, i64],
    "wf2cu_sce_backward": [vp, i64, vp, i64, vp, i64, vp, i64, i64, i64, i32],
    "wf2cu_sgd_step": [vp, i64, vp, i64, i64, f32],
    "wf2cu_adam_step": [vp, i64, vp, i64, vp, i64, vp, i64, i64, f32, vp, f32, f32, f32, f32],
    "wf2cu_embedding_gather": [vp, i64, i64, vp, vp, i64, vp],
    "wf2cu_layernorm": [vp, i64, vp, i64, vp, i64, vp, i64, i64, i64, f32],
    "wf2cu_attn_heads_softmax": [vp, i64, i64, i64, i64, i64, f32, vp, i64],
    "wf2cu_attn_pv": [vp, i64, vp, i64, i64, i64, i64, i64, vp, i64],
    "wf2cu_attn_fused": [vp, i64, i64, i64, i64, i64, f32, vp, i64],
    "wf2cu_im2col": [vp, i64, vp, i64, GP], "wf2cu_col2im": [vp, i64, vp, i64, GP],
    "wf2cu_unfold": [VP, vp, i64, GP], "wf2cu_fold": [vp, i64, VP, GP, i32],
    "wf2cu_pool2d": [i32, VP, VP, GP], "wf2cu_pool2d_backward": [i32, VP, VP, VP, VP, GP, i32, i32],
    "wf2cu_comm_unique_id": [vp], "wf2cu_comm_init": [i32, i32, vp],
    "wf2cu_allreduce_sum_f32": [vp, i64, i64], "wf2cu_comm_destroy": [],
    "wf2cu_p2p_export": [vp, vp], "wf2cu_p2p_signal_export": [vp], "wf2cu_p2p_attach": [i32, i32, vp, vp, vp, vp, vp],
    "wf2cu_p2p_allreduce_adam_step": [vp, i64, i64, vp, i64, vp, i64, i64, f32, vp, f32, f32, f32],
    "wf2cu_p2p_status": [C.POINTER(i32)], "wf2cu_p2p_detach": [],
}

# enums of include/waffe2cuda.h
ADD, SUB, MUL, DIV, COPY = range(5)
UNARY = {n: i for i, n in enumerate(
    ["sin", "cos", "tan", "asin", "acos", "atan", "sinh", "cosh", "tanh", "abs", "exp", "loge", "log2", "log10",
     "log1p", "sqrt", "cbrt", "sign", "square", "neg", "recip"])}
CMP = {"eq": 0, "lt": 1, "le": 2, "gt": 3, "ge": 4, "=": 0, "<": 1, "<=": 2, ">": 3, ">=": 4}
P_LOAD, P_STORE, P_IMM, P_ADD, P_SUB, P_MUL, P_DIV, P_UNARY, P_POW, P_CMPSEL, P_MOV = range(11)

_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library and attach prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python cl-waffe2_b200/build.py` "
                          "(there is no CPU fallback for the CUDATensor backend)")
    lib = C.CDLL(LIB_PATH)
    for name, args in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.argtypes = args
        fn.restype = C.c_int
    lib.wf2cu_last_error.argtypes = []
    lib.wf2cu_last_error.restype = C.c_char_p
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise Wf2CudaError(rc, load().wf2cu_last_error().decode("utf-8", "replace"))


_inited_device = None


def init(device: int = 0) -> None:
    """wf2cu_init; raises Wf2CudaError when no sm_100 device is visible."""
    global _inited_device
    if _inited_device == device:
        return
    check(load().wf2cu_init(device))
    _inited_device = device


def shutdown() -> None:
    global _inited_device
    if _lib is not None:
        _lib.wf2cu_shutdown()
    _inited_device = None


def make_view(base: int, offset: int, shape, stride) -> View:
    v = View()
    n = len(shape)
    if n > MAX_RANK:
        raise Wf2CudaError(2, f"rank {n} exceeds WF2CU_MAX_RANK={MAX_RANK}")
    v.base = base
    v.offset = int(offset)
    v.rank = n
    for i in range(n):
        v.shape[i] = int(shape[i])
        v.stride[i] = int(stride[i])
    return v
