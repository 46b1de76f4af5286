"""oracle/gpt2.py — TEST INFRASTRUCTURE (not product): numpy restatement of the reference's GPT-2
example forward (examples/gpt-2/model.lisp, utils.lisp), used as the checker for
`cl-waffe2_b200/gpt2.py`.  Parity UNPINNED: the reference holds no golden vectors or tests for this
example (it needs downloaded GPT-2 weights) and no Common Lisp exists here to run it; the
restatement follows the source line by line, each function citing the lines it follows.

`dtype=np.float64` gives the ground truth, `np.float32` a reference-precision run."""
from __future__ import annotations

import numpy as np


def layer_norm(x, g, b, eps=1e-5):
    """LayerNorm-Revisit, examples/gpt-2/utils.lisp:114-127 (same formula as source/nn/norms.lisp:72-82)."""
    u = x.mean(axis=-1, keepdims=True)
    s = ((x - u) ** 2).mean(axis=-1, keepdims=True)
    x = (x - u) / np.sqrt(s + x.dtype.type(eps))
    return x * g + b


def gelu_tanh(x):
    """!gelu-lisptanh, utils.lisp:129-137."""
    t = x.dtype.type
    c = t(np.float32(np.sqrt(2.0 / np.pi)))     # (coerce (sqrt (/ 2.0 pi)) 'single-float)
    return t(0.5) * x * (t(1) + np.tanh(c * (x + t(0.044715) * x ** 3)))


def split_heads(x, which, n_head):
    """utils.lisp:204-215: slice the last axis, reshape to heads, permute; K -> (B H d N)."""
    E = x.shape[-1] // 3
    sl = {"Q": slice(0, E), "K": slice(E, 2 * E), "V": slice(2 * E, 3 * E)}[which]
    y = x[..., sl]
    B, N, _ = y.shape
    y = y.reshape(B, N, n_head, E // n_head)
    return y.transpose(0, 2, 3, 1) if which == "K" else y.transpose(0, 2, 1, 3)


def merge_heads(x):
    """utils.lisp:217-220."""
    B, H, N, d = x.shape
    return x.transpose(0, 2, 1, 3).reshape(B, N, H * d)


def softmax_no_shift(x, axis=1):
    """(!softmax x :avoid-overflow nil), source/nn/activation.lisp:126-159: exp(x) / sum(exp(x), axis)."""
    e = np.exp(x)
    return e / e.sum(axis=axis, keepdims=True)


def self_attention(x, n_head):
    """SelfAttention without a past, utils.lisp:222-244 — as written: scale 1/N (N = last axis of w),
    softmax over axis 1 (the heads axis of the 4-D tensor), no causal mask."""
    K = split_heads(x, "K", n_head)
    V = split_heads(x, "V", n_head)
    Q = split_heads(x, "Q", n_head)
    w = Q @ K
    w = w / w.dtype.type(w.shape[-1])
    w = softmax_no_shift(w, axis=1)
    return merge_heads(w @ V)


def layer(x, W, i, n_head):
    """gpt2layer-call, model.lisp:93-137."""
    p = f"h{i}."
    a = layer_norm(x, W[p + "ln_1.g"], W[p + "ln_1.b"])
    a = a @ W[p + "attn.c_attn.w"] + W[p + "attn.c_attn.b"]
    a = self_attention(a, n_head)
    a = a @ W[p + "attn.c_proj.w"] + W[p + "attn.c_proj.b"]
    x = x + a
    m = layer_norm(x, W[p + "ln_2.g"], W[p + "ln_2.b"])
    m = m @ W[p + "mlp.c_fc.w"] + W[p + "mlp.c_fc.b"]
    m = gelu_tanh(m)
    m = m @ W[p + "mlp.c_proj.w"] + W[p + "mlp.c_proj.b"]
    return x + m


def embed(ctx, wte, wpe):
    """GPT2PositionalEmbedding, utils.lisp:19-112: wte[round(ctx[b,p])] + wpe[p]; ids are floats."""
    ids = np.rint(ctx).astype(np.int64)
    return wte[ids] + wpe[np.arange(ctx.shape[1])][None, :, :]


def forward(ctx, weights, cfg, dtype=np.float64, lm_head=False):
    """(call (GPT2) prev), model.lisp:162-192."""
    W = {k: v.astype(dtype) for k, v in weights.items()}
    x = embed(np.asarray(ctx), W["wte"], W["wpe"])
    for i in range(cfg["n_layer"]):
        x = layer(x, W, i, cfg["n_head"])
    x = layer_norm(x, W["ln_f.g"], W["ln_f.b"])
    if lm_head:
        x = x @ W["wte"].T
    return x
