"""GPT-2 forward on CUDATensor (BASELINE.json config 5: examples/gpt-2, GPT-2 small, random-init
weights, seq 1024, batch 8), composed from the backend's nodes exactly as the reference example
composes them — including its quirks, reproduced as written:

  * `SelfAttention` (examples/gpt-2/utils.lisp:222-244): w = Q.K ; w / N where N = the LAST axis
    of w (sequence length, not sqrt(head_dim)) ; `(!softmax w :avoid-overflow nil)` with the
    default `:axis 1`, i.e. normalised over the HEADS axis of the (B H N N) tensor ; no causal mask.
  * `split-heads` / `merge-heads` (utils.lisp:204-220): slice of the fused QKV projection,
    reshape to heads, permute; K is permuted to (B H d N).
  * `LayerNorm-Revisit` (utils.lisp:114-127) and nn `LayerNorm` (source/nn/norms.lisp:44-85):
    u = mean(x,-1); s = mean((x-u)^2,-1); (x-u)/sqrt(s+eps) * g + b, eps = 1e-5.
  * `!gelu-lisptanh` (utils.lisp:129-137): 0.5 x (1 + tanh(sqrt(2/pi)(x + 0.044715 x^3))).
  * `GPT2PositionalEmbedding` (utils.lisp:19-112): out[b,p,:] = wte[round(ctx[b,p])] + wpe[p];
    token ids are floats.
  * `lm-head` (model.lisp:191-192): x . wte^T.

Each step below is one call of the C ABI (one kernel, or split+GEMM for `!matmul`)."""
from __future__ import annotations

import os

from typing import Dict, List, Optional

import numpy as np

from . import _lib, ops
from .tensor import CUDATensor, as_strided, from_numpy, make_tensor, reshape_contiguous, view, T

CONFIG_SMALL = dict(n_vocab=50257, n_ctx=1024, n_emb=768, n_head=12, n_layer=12)


def init_weights(cfg: Dict[str, int], seed: int = 0, std: float = 0.02) -> Dict[str, np.ndarray]:
    """Random-init weights with the shapes the example loads from .npy (model.lisp:64-86,145-160):
    N(0, 0.02) matrices, LayerNorm gains 1 + N(0,0.02), biases N(0, 0.02).  Host arrays, so the same
    numbers feed the oracle and the device."""
    rng = np.random.default_rng(seed)
    E, V, C, L = cfg["n_emb"], cfg["n_vocab"], cfg["n_ctx"], cfg["n_layer"]
    f = lambda *s: rng.normal(0.0, std, size=s).astype(np.float32)
    w = {"wte": f(V, E), "wpe": f(C, E), "ln_f.g": (1 + f(E)).astype(np.float32), "ln_f.b": f(E)}
    for i in range(L):
        p = f"h{i}."
        w[p + "ln_1.g"] = (1 + f(E)).astype(np.float32); w[p + "ln_1.b"] = f(E)
        w[p + "ln_2.g"] = (1 + f(E)).astype(np.float32); w[p + "ln_2.b"] = f(E)
        w[p + "attn.c_attn.w"] = f(E, 3 * E); w[p + "attn.c_attn.b"] = f(3 * E)
        w[p + "attn.c_proj.w"] = f(E, E); w[p + "attn.c_proj.b"] = f(E)
        w[p + "mlp.c_fc.w"] = f(E, 4 * E); w[p + "mlp.c_fc.b"] = f(4 * E)
        w[p + "mlp.c_proj.w"] = f(4 * E, E); w[p + "mlp.c_proj.b"] = f(E)
    return w


FUSE_LAYERNORM = True


def _bcast_last(vec1d: CUDATensor, like: CUDATensor) -> CUDATensor:
    """`(%transform b[i] -> [~ i])`: a (E) vector broadcast over every leading axis of `like`."""
    r = like.dims()
    t = reshape_contiguous(vec1d, [1] * (r - 1) + [like.shape[-1]])
    return view(t, *([[":broadcast", n] for n in like.shape[:-1]] + [T]))


def _bcast_keep(stat: CUDATensor, like: CUDATensor) -> CUDATensor:
    """A keepdims statistic (... 1) broadcast along the last axis of `like`."""
    return view(stat, *([T] * (like.dims() - 1) + [[":broadcast", like.shape[-1]]]))


EMIT_PLANES = True      # producers write the operand plane of the MatmulNode that follows (wf2cu_emit_operand_planes)
EMIT_FROM_GEMM = os.environ.get("WF2CU_EMIT_FROM_GEMM", "1") != "0"   # ... also the gelu projection's epilogue (A/B switch)


def layer_norm(x: CUDATensor, g: CUDATensor, b: CUDATensor, eps: float = 1e-5, feeds_matmul: bool = False) -> CUDATensor:
    """LayerNorm-Revisit (examples/gpt-2/utils.lisp:114-127): one kernel when FUSE_LAYERNORM (the ten-node chain
    with the same per-element operations, wf2cu_layernorm), node by node otherwise.  feeds_matmul: the result is the A operand
    of the projection issued next, so the kernel also writes its tensor-core operand plane (same bits as the product's pre-pass)."""
    shape = x.shape
    if FUSE_LAYERNORM and x.is_contiguous() and not x.permuted_p() and shape[-1] % 4 == 0 and shape[-1] <= 1024:
        out = make_tensor(shape)
        rows = x.numel() // shape[-1]
        if feeds_matmul and EMIT_PLANES:
            _lib.check(_lib.load().wf2cu_emit_operand_planes(1))
        _lib.check(_lib.load().wf2cu_layernorm(x.vec.ptr, x.lower().offset, g.vec.ptr, g.lower().offset, b.vec.ptr, b.lower().offset,
                                               out.vec.ptr, out.lower().offset, rows, shape[-1], float(eps)))
        return out
    u = make_tensor(shape[:-1] + [1])
    ops.mean(x, axis=-1, out=u)
    d = ops.sub(x, _bcast_keep(u, x))                       # x - u
    sq = ops.expt(d, 2)                                     # (!expt (!sub x u) 2)
    s = make_tensor(shape[:-1] + [1])
    ops.mean(sq, axis=-1, out=s)
    ops.add_(s, eps)                                        # s + eps
    ops.unary("sqrt", s, out=s)
    ops.div_(d, _bcast_keep(s, x))                          # (x - u) / sqrt(s + eps)
    ops.mul_(d, _bcast_last(g, x))
    ops.add_(d, _bcast_last(b, x))
    return d


def split_heads(qkv: CUDATensor, which: str, n_head: int) -> CUDATensor:
    """utils.lisp:204-215.  qkv: dense (B N 3E).  Q, V -> (B H N d); K -> (B H d N).  Views only."""
    B, N, E3 = qkv.shape
    E = E3 // 3
    d = E // n_head
    off = {"Q": 0, "K": E, "V": 2 * E}[which]
    if which == "K":
        return as_strided(qkv, [B, n_head, d, N], [N * E3, d, 1, E3], off)
    return as_strided(qkv, [B, n_head, N, d], [N * E3, d, E3, 1], off)


def merge_heads(x: CUDATensor) -> CUDATensor:
    """utils.lisp:217-220: (B H N d) -> permute (B N H d) -> reshape (B N E) (one transpose copy)."""
    B, H, N, d = x.shape
    xc = ops.contiguous(x)
    p = as_strided(xc, [B, N, H, d], [H * N * d, d, N * d, 1], 0)
    return reshape_contiguous(ops.copy(p), [B, N, H * d])


FUSE_ATTENTION = True
ONE_CALL_ATTENTION = True   # wf2cu_attn_fused: one pass over the heads, the weights never leave the library's workspace normalised


def self_attention(qkv: CUDATensor, n_head: int) -> CUDATensor:
    """SelfAttention (utils.lisp:222-244) without a past: as written in the example.
    FUSE_ATTENTION: two launches (wf2cu_attn_heads_softmax, wf2cu_attn_pv: csrc/attention.cu) read Q, K and V straight from
    the QKV activation, write the (B H N N) weights once and leave the result in merged-heads layout; shapes the
    library does not take (WF2CU_ERR_UNSUPPORTED) run node by node below, exactly as the example composes them."""
    if FUSE_ATTENTION and qkv.dims() == 3 and qkv.is_contiguous() and not qkv.permuted_p():
        lib = _lib.load()
        B, N, E3 = qkv.shape
        E = E3 // 3
        off = qkv.lower().offset
        if ONE_CALL_ATTENTION:
            out = make_tensor([B, N, E])
            if EMIT_PLANES:                                  # the output projection is issued next
                _lib.check(lib.wf2cu_emit_operand_planes(1))
            rc = lib.wf2cu_attn_fused(qkv.vec.ptr, off, B, N, n_head, E // n_head, float(N), out.vec.ptr, out.lower().offset)
            if rc == 0:
                return out
            if rc != 3:
                _lib.check(rc)
            _lib.check(lib.wf2cu_emit_operand_planes(0))
        w = make_tensor([B, n_head, N, N])
        rc = lib.wf2cu_attn_heads_softmax(qkv.vec.ptr, off, B, N, n_head, E // n_head, float(N), w.vec.ptr, w.lower().offset)
        if rc == 0:
            out = make_tensor([B, N, E])
            if EMIT_PLANES:                                  # the output projection is issued next
                _lib.check(lib.wf2cu_emit_operand_planes(1))
            _lib.check(lib.wf2cu_attn_pv(w.vec.ptr, w.lower().offset, qkv.vec.ptr, off, B, N, n_head, E // n_head, out.vec.ptr, out.lower().offset))
            return out
        if rc != 3:                              # anything but WF2CU_ERR_UNSUPPORTED is an error
            _lib.check(rc)
    Q = split_heads(qkv, "Q", n_head)
    K = split_heads(qkv, "K", n_head)
    V = split_heads(qkv, "V", n_head)
    w = ops.matmul(Q, K)                                    # (B H N N); operands made contiguous (adjust-layout)
    # (!div w (make-tensor N)) then (!softmax w :avoid-overflow nil), default axis=1 -> over heads; one pass
    ops.softmax(w, avoid_overflow=False, axis=1, out=w, pre_divisor=float(w.shape[-1]))
    a = ops.matmul(w, V)                                    # (B H N d)
    return merge_heads(a)


FUSE_EPILOGUES = True


def _linear(x: CUDATensor, w: CUDATensor, b: CUDATensor, gelu: bool = False, residual: Optional[CUDATensor] = None,
            feeds_matmul: bool = False) -> CUDATensor:
    """`(!matmul x w)` + `(!add (%transform b[i] -> [~ i]))` (model.lisp:118-120), optionally followed by the node the block
    puts next: `!gelu` (after the first MLP projection) or the residual `(!add x ...)` (after both output projections).
    feeds_matmul: the result is the left operand of the `!matmul` issued next; the epilogue that applies bias / gelu also writes
    that product's operand plane (wf2cu_emit_operand_planes, as layer_norm does)."""
    if x.dims() == 3 and x.is_contiguous() and not x.permuted_p():
        # W is shared by the batch, so (B N in) . (in out) is one (B*N in) . (in out) product; the bias add rides in the
        # GEMM epilogue (wf2cu_gemm_ex — the MatMul + Add fusion of lisp/fusion.lisp), same arithmetic, one pass less;
        # FUSE_EPILOGUES: the gelu / residual node too (wf2cu_gemm_ex2), bit-identical to the separate nodes
        B_, N_, _ = x.shape
        x2 = reshape_contiguous(x, [B_ * N_, x.shape[2]])
        fuse = FUSE_EPILOGUES and (gelu or residual is not None) and (residual is None or (residual.is_contiguous() and not residual.permuted_p()))
        if fuse:
            r2 = reshape_contiguous(residual, [B_ * N_, residual.shape[2]]) if residual is not None else None
            if feeds_matmul and EMIT_PLANES and EMIT_FROM_GEMM:
                _lib.check(_lib.load().wf2cu_emit_operand_planes(1))
            y = ops.matmul(x2, w, bias=b, gelu=gelu, residual=r2)
            return reshape_contiguous(y, [B_, N_, y.shape[1]])
        y = reshape_contiguous(ops.matmul(x2, w, bias=b), [B_, N_, w.shape[1]])
    else:
        y = ops.matmul(x, w)                                # (B N in) . (in out): W shared by the batch
        ops.add_(y, _bcast_last(b, y))
    if gelu:
        ops.gelu(y, out=y)
    if residual is not None:
        y = ops.add(residual, y)
    return y


class GPT2Layer:
    """model.lisp:47-137."""

    def __init__(self, weights: Dict[str, np.ndarray], i: int, n_head: int):
        p = f"h{i}."
        g = lambda k: from_numpy(weights[p + k])
        self.n_head = n_head
        self.ln_1_g, self.ln_1_b, self.ln_2_g, self.ln_2_b = g("ln_1.g"), g("ln_1.b"), g("ln_2.g"), g("ln_2.b")
        self.attn_w, self.attn_b = g("attn.c_attn.w"), g("attn.c_attn.b")
        self.proj_w, self.proj_b = g("attn.c_proj.w"), g("attn.c_proj.b")
        self.fc_w, self.fc_b = g("mlp.c_fc.w"), g("mlp.c_fc.b")
        self.fc_proj_w, self.fc_proj_b = g("mlp.c_proj.w"), g("mlp.c_proj.b")

    def __call__(self, x: CUDATensor) -> CUDATensor:
        h = layer_norm(x, self.ln_1_g, self.ln_1_b, feeds_matmul=True)
        qkv = _linear(h, self.attn_w, self.attn_b)
        a = self_attention(qkv, self.n_head)
        x = _linear(a, self.proj_w, self.proj_b, residual=x)            # (!add x (proj a)): residual
        h = layer_norm(x, self.ln_2_g, self.ln_2_b, feeds_matmul=True)
        m = _linear(h, self.fc_w, self.fc_b, gelu=True, feeds_matmul=True)   # (!gelu (fc h)), the left operand of the projection below
        return _linear(m, self.fc_proj_w, self.fc_proj_b, residual=x)   # (!add x (fc-proj m)): residual


class GPT2:
    """model.lisp:139-192: embedding -> n_layer blocks -> LayerNorm [-> lm-head]."""

    def __init__(self, cfg: Dict[str, int] = None, weights: Optional[Dict[str, np.ndarray]] = None, seed: int = 0,
                 constant_weights: bool = True):
        self.cfg = dict(CONFIG_SMALL if cfg is None else cfg)
        w = weights if weights is not None else init_weights(self.cfg, seed)
        self.wte, self.wpe = from_numpy(w["wte"]), from_numpy(w["wpe"])
        self.ln_g, self.ln_b = from_numpy(w["ln_f.g"]), from_numpy(w["ln_f.b"])
        self.layers: List[GPT2Layer] = [GPT2Layer(w, i, self.cfg["n_head"]) for i in range(self.cfg["n_layer"])]
        if constant_weights:
            self.declare_constant(True)

    def declare_constant(self, on: bool = True) -> None:
        """The example only runs the model forward (gpt2-model.lisp loads the weights once): the weight matrices are declared
        constant operands (wf2cu_gemm_const_operand), so MatmulNode keeps their tensor-core operand planes between calls
        instead of re-splitting 48 matrices per forward."""
        lib = _lib.load()
        mats = [self.wte] + [m for l in self.layers for m in (l.attn_w, l.proj_w, l.fc_w, l.fc_proj_w)]
        for t in mats:
            _lib.check(lib.wf2cu_gemm_const_operand(t.vec.ptr, 0, int(np.prod(t.shape)), 1 if on else 0))

    def embed(self, ctx: CUDATensor) -> CUDATensor:
        """GPT2PositionalEmbedding (utils.lisp:19-112); ctx: (B N) float token ids."""
        B, N = ctx.shape
        E = self.cfg["n_emb"]
        out = make_tensor([B, N, E])
        lc = ctx.lower()
        _lib.check(_lib.load().wf2cu_embedding_gather(ctx.vec.ptr + 4 * lc.offset, B, N, self.wte.vec.ptr,
                                                      self.wpe.vec.ptr, E, out.vec.ptr))
        return out

    def forward(self, ctx: CUDATensor, lm_head: bool = False) -> CUDATensor:
        x = self.embed(ctx)
        for layer in self.layers:
            x = layer(x)
        x = layer_norm(x, self.ln_g, self.ln_b)
        if lm_head:
            x = ops.matmul(x, ops.t(self.wte))              # (B N E) . (V E)^T
        return x

    __call__ = forward
