"""Strict-precision ("fp32 mode") autograd glue: the same modules, reference-level numerics.

The reference is fp32 end to end (default-dtype nn modules in models/vit.py run unchanged by
going_modular/engine.py:50-66). BASELINE.json asks for logits within 1e-3 relative of it in an "fp32 mode"; the
throughput path (b200vit.functional, bf16 GEMM operands) lands at ~4e-3. Under `b200vit.precision("fp32")` the
modules dispatch here instead:

  * activations, gradients and everything elementwise stay fp32 between kernels;
  * every contraction is still ONE launch of the tcgen05 GEMM kernel, fed split operands concatenated along the
    contraction dimension, A' = [hi | lo | hi], B' = [hi | hi | lo], K' = 3K (csrc/fp32_mode.cu): the fp32
    accumulator in tensor memory receives a_hi b_hi + a_lo b_hi + a_hi b_lo, about 2^-16 relative per product;
  * attention, GELU and bias gradients run in fp32 CUDA-core kernels.

It is 3-4x slower than the bf16 path and exists to pin numerics (tests/test_fp32_mode_gpu.py), not to be fast.
Orchestration only: torch allocates, every FLOP runs in libb200vit.so.
"""
from __future__ import annotations

import math
import weakref
from typing import Optional

import torch

from . import kernels as K
from .functional import (HEAD_DIM, LN_EPS, _empty, _f32c, _require_cuda, _zeros, next_dropout_key)

PAT_A, PAT_B = 0, 1     # (hi, lo, hi) / (hi, hi, lo)


def _split(src: torch.Tensor, pattern: int, concat: str, p: float = 0.0, key=(0, 0), row_period: int = 0,
           rows: Optional[int] = None) -> torch.Tensor:
    """bf16x3 split of a dense f32 matrix [R, C] -> [R, 3C] (concat='cols') or [3R, C] (concat='rows')."""
    r_src, c = src.shape
    r = r_src if rows is None else rows
    if concat == "cols":
        dst = torch.empty((r, 3 * c), dtype=torch.bfloat16, device=src.device)
        K.split_bf16x3(src, c, r, c, dst, 3 * c, c, pattern, p, key[0], key[1], row_period)
    else:
        dst = torch.empty((3 * r, c), dtype=torch.bfloat16, device=src.device)
        K.split_bf16x3(src, c, r, c, dst, c, r * c, pattern, p, key[0], key[1], row_period)
    return dst


# split copies of the weights, re-made only when the master changed (same scheme as functional.bf16_shadow)
_WSPLIT: dict = {}


def weight_split(p: torch.Tensor, concat: str) -> torch.Tensor:
    """B-side split of a weight [N, K] (conv weights are viewed as [D, C*P*P]): 'cols' -> [N, 3K] for the forward
    NT product, 'rows' -> [3N, K] for the input-gradient NN product."""
    key = (id(p), concat)
    e = _WSPLIT.get(key)
    if e is not None and e[3]() is p and e[1] == p._version and e[2] == p.data_ptr() and e[0].device == p.device:
        return e[0]
    w2d = p.detach().reshape(p.shape[0], -1)
    if not w2d.is_contiguous():
        w2d = w2d.contiguous()
    t = _split(w2d, PAT_B, concat)
    _WSPLIT[key] = (t, p._version, p.data_ptr(), weakref.ref(p, lambda _r, k=key: _WSPLIT.pop(k, None)))
    return t


def _lin_fwd(x2d, w, b, residual=None, p=0.0, key=(0, 0)):
    """y = dropout(x . W^T + b) + residual   (nn.Linear + nn.Dropout + the `+ x` of models/vit.py:167-168)."""
    r, k = x2d.shape
    n = w.shape[0]
    xa = _split(x2d, PAT_A, "cols")
    wb = weight_split(w, "cols")
    y = _empty((r, n), torch.float32, x2d)
    if residual is None and p == 0.0:
        K.gemm(K.OP_NT, K.EPI_F32, r, n, 3 * k, xa, 3 * k, wb, 3 * k, y, n, bias=b)
    else:
        K.gemm(K.OP_NT, K.EPI_BIAS_RES, r, n, 3 * k, xa, 3 * k, wb, 3 * k, y, n, bias=b, aux=residual, ldaux=n,
               dropout_p=p, seed=key[0], offset=key[1])
    return y


def _lin_bwd(g, x2d, w, p=0.0, key=(0, 0), need_dx=True):
    """Gradients of y = dropout(x . W^T + b): (dx, dW, db) from g = dL/dy (the mask is applied while g is read)."""
    r, n = g.shape
    k = x2d.shape[1]
    db = _empty((n,), torch.float32, g)
    K.colsum_f32(g, n, r, n, db, p, key[0], key[1])
    ga_r = _split(g, PAT_A, "rows", p, key)
    xb_r = _split(x2d, PAT_B, "rows")
    dw = _zeros((n, k), g)
    K.gemm(K.OP_TN, K.EPI_ACCUM, n, k, 3 * r, ga_r, n, xb_r, k, dw, k, split_k=0)
    dx = None
    if need_dx:
        ga_c = _split(g, PAT_A, "cols", p, key)
        wb_r = weight_split(w, "rows")
        dx = _empty((r, k), torch.float32, g)
        K.gemm(K.OP_NN, K.EPI_F32, r, k, 3 * n, ga_c, 3 * n, wb_r, k, dx, k)
    return dx, dw, db


def _ln_fwd(x2d, w, b, eps):
    r, d = x2d.shape
    y = _empty((r, d), torch.float32, x2d)
    mean = _empty((r,), torch.float32, x2d)
    rstd = _empty((r,), torch.float32, x2d)
    K.layernorm_fwd(x2d, d, w, b, None, y, mean, rstd, r, d, eps)
    return y, mean, rstd


def _ln_bwd(dy, x2d, mean, rstd, w, gin):
    r, d = x2d.shape
    gx = _empty((r, d), torch.float32, x2d)
    dgamma = _empty((d,), torch.float32, x2d)
    dbeta = _empty((d,), torch.float32, x2d)
    K.layernorm_bwd(None, dy, x2d, d, mean, rstd, w, gin, gx, d, None, dgamma, dbeta, None, r, d)
    return gx, dgamma, dbeta


def _msa_fwd(x2d, bn, t, heads, ln_w, ln_b, w_in, b_in, w_out, b_out, residual, eps):
    r, d = x2d.shape
    if d != heads * HEAD_DIM:
        raise NotImplementedError(f"b200vit attention kernels need head_dim == 64 (embedding_dim {d}, heads {heads})")
    ln, mean, rstd = _ln_fwd(x2d, ln_w, ln_b, eps)
    qkv = _lin_fwd(ln, w_in, b_in)
    o = _empty((r, d), torch.float32, x2d)
    lse = _empty((bn, heads, t), torch.float32, x2d)
    K.attention_f32_fwd(qkv, o, lse, bn, t, heads, HEAD_DIM, 1.0 / math.sqrt(HEAD_DIM))
    y = _lin_fwd(o, w_out, b_out, residual=residual)
    return y, (x2d, mean, rstd, ln, qkv, o, lse)


def _msa_bwd(g, saved, bn, t, heads, ln_w, w_in, w_out, has_residual):
    x2d, mean, rstd, ln, qkv, o, lse = saved
    do, dw_out, db_out = _lin_bwd(g, o, w_out)
    delta = _empty((bn, heads, t), torch.float32, x2d)
    dqkv = torch.empty_like(qkv)
    K.attention_f32_bwd(qkv, o, do, lse, delta, dqkv, bn, t, heads, HEAD_DIM, 1.0 / math.sqrt(HEAD_DIM))
    dln, dw_in, db_in = _lin_bwd(dqkv, ln, w_in)
    gx, dgamma, dbeta = _ln_bwd(dln, x2d, mean, rstd, ln_w, g if has_residual else None)
    return gx, (dgamma, dbeta, dw_in, db_in, dw_out, db_out)


def _mlp_fwd(x2d, ln_w, ln_b, w1, b1, w2, b2, p, residual, eps):
    ln, mean, rstd = _ln_fwd(x2d, ln_w, ln_b, eps)
    key1 = next_dropout_key(x2d.device) if p > 0 else (0, 0)
    key2 = next_dropout_key(x2d.device) if p > 0 else (0, 0)
    a = _lin_fwd(ln, w1, b1)
    h = torch.empty_like(a)
    K.gelu_f32_fwd(a, h, p, key1[0], key1[1])
    y = _lin_fwd(h, w2, b2, residual=residual, p=p, key=key2)
    return y, (x2d, mean, rstd, ln, a, h), (p, key1, key2)


def _mlp_bwd(g, saved, drop, ln_w, w1, w2, has_residual):
    x2d, mean, rstd, ln, a, h = saved
    p, key1, key2 = drop
    dh, dw2, db2 = _lin_bwd(g, h, w2, p, key2)
    da = torch.empty_like(a)
    K.gelu_f32_bwd(a, dh, da, p, key1[0], key1[1])
    dln, dw1, db1 = _lin_bwd(da, ln, w1)
    gx, dgamma, dbeta = _ln_bwd(dln, x2d, mean, rstd, ln_w, g if has_residual else None)
    return gx, (dgamma, dbeta, dw1, db1, dw2, db2)


def _no_attn_dropout(attn_p: float, training: bool) -> None:
    if training and attn_p > 0:
        raise NotImplementedError("fp32 mode does not implement attn_dropout > 0 (the reference default is 0); "
                                  "use the default bf16 precision for that option")


class PatchEmbedF32Fn(torch.autograd.Function):
    """PatchEmbedding.forward (models/vit.py:45-67) at fp32-equivalent accuracy."""

    @staticmethod
    def forward(ctx, img, w, b, cls, pos, patch: int, p_drop: float, training: bool):
        _require_cuda(img, "PatchEmbedding")
        if img.dtype == torch.uint8:        # the reference's ToDtype(float32, scale=True) input transform
            img = img.to(torch.float32).div_(255.0)
        img = _f32c(img)
        bn, c, h, wd = img.shape
        d = w.shape[0]
        np_ = (h // patch) * (wd // patch)
        t = np_ + 1
        if pos.shape[1] != t:
            raise RuntimeError(f"image of {h}x{wd} gives {t} tokens but position_embedding has {pos.shape[1]}")
        kdim = c * patch * patch
        patches = K.im2col_patches_f32(img, patch)
        pa = _split(patches, PAT_A, "cols")
        wb = weight_split(w, "cols")
        x = _empty((bn, t, d), torch.float32, img)
        p = float(p_drop) if training else 0.0
        seed, off = next_dropout_key(img.device) if p > 0 else (0, 0)
        K.gemm(K.OP_NT, K.EPI_PATCH, bn * np_, d, 3 * kdim, pa, 3 * kdim, wb, 3 * kdim, x, d, bias=b, aux=pos,
               ldaux=d, dropout_p=p, seed=seed, offset=off, row_period=np_)
        K.class_token_rows(cls, pos, x, bn, t, d, p, seed, off)
        ctx.save_for_backward(patches)
        ctx.meta = (bn, t, d, kdim, p, seed, off, w.shape)
        return x

    @staticmethod
    def backward(ctx, gx):
        (patches,) = ctx.saved_tensors
        bn, t, d, kdim, p, seed, off, wshape = ctx.meta
        gx = _f32c(gx)
        g_cls = _empty((1, 1, d), torch.float32, gx)
        g_pos = _empty((1, t, d), torch.float32, gx)
        g_b = _empty((d,), torch.float32, gx)
        K.embed_param_grads(gx, g_cls, g_pos, g_b, bn, t, d, p, seed, off)
        rows = bn * (t - 1)
        # the patch rows of gx, dropout mask applied, class rows skipped -- read in place by the split kernel
        ga = _split(gx.view(bn * t, d), PAT_A, "rows", p, (seed, off), row_period=t - 1, rows=rows)
        pb = _split(patches, PAT_B, "rows")
        gw = _zeros((d, kdim), gx)
        K.gemm(K.OP_TN, K.EPI_ACCUM, d, kdim, 3 * rows, ga, d, pb, kdim, gw, kdim, split_k=0)
        return None, gw.view(wshape), g_b, g_cls, g_pos, None, None, None


class MSAF32Fn(torch.autograd.Function):
    """MultiHeadSelfAttentionBlock.forward (models/vit.py:92-98)."""

    @staticmethod
    def forward(ctx, x, ln_w, ln_b, w_in, b_in, w_out, b_out, heads: int, attn_p: float, training: bool, eps: float):
        _require_cuda(x, "MultiHeadSelfAttentionBlock")
        _no_attn_dropout(attn_p, training)
        x = _f32c(x)
        bn, t, d = x.shape
        y, saved = _msa_fwd(x.view(bn * t, d), bn, t, heads, ln_w, ln_b, w_in, b_in, w_out, b_out, None, eps)
        ctx.save_for_backward(ln_w, w_in, w_out, *saved)
        ctx.meta = (bn, t, heads)
        return y.view(bn, t, d)

    @staticmethod
    def backward(ctx, g):
        ln_w, w_in, w_out, *saved = ctx.saved_tensors
        bn, t, heads = ctx.meta
        d = g.shape[-1]
        gx, (dgam, dbet, dw_in, db_in, dw_out, db_out) = _msa_bwd(_f32c(g).view(bn * t, d), tuple(saved), bn, t, heads,
                                                                  ln_w, w_in, w_out, has_residual=False)
        return gx.view(bn, t, d), dgam, dbet, dw_in, db_in, dw_out, db_out, None, None, None, None


class MLPF32Fn(torch.autograd.Function):
    """MLPBlock.forward (models/vit.py:128-131)."""

    @staticmethod
    def forward(ctx, x, ln_w, ln_b, w1, b1, w2, b2, p: float, training: bool, eps: float):
        _require_cuda(x, "MLPBlock")
        x = _f32c(x)
        shape = x.shape
        d = shape[-1]
        y, saved, drop = _mlp_fwd(x.view(-1, d), ln_w, ln_b, w1, b1, w2, b2, float(p) if training else 0.0, None, eps)
        ctx.save_for_backward(ln_w, w1, w2, *saved)
        ctx.drop = drop
        return y.view(shape)

    @staticmethod
    def backward(ctx, g):
        ln_w, w1, w2, *saved = ctx.saved_tensors
        shape = g.shape
        d = shape[-1]
        gx, (dgam, dbet, dw1, db1, dw2, db2) = _mlp_bwd(_f32c(g).view(-1, d), tuple(saved), ctx.drop, ln_w, w1, w2,
                                                        has_residual=False)
        return gx.view(shape), dgam, dbet, dw1, db1, dw2, db2, None, None, None


class EncoderBlockF32Fn(torch.autograd.Function):
    """TransformerEncoderBlock.forward (models/vit.py:166-169): x1 = msa(x) + x ; x2 = mlp(x1) + x1."""

    @staticmethod
    def forward(ctx, x, a_ln_w, a_ln_b, w_in, b_in, w_out, b_out, m_ln_w, m_ln_b, w1, b1, w2, b2, heads: int, p: float,
                training: bool, attn_p: float, eps_msa: float, eps_mlp: float):
        _require_cuda(x, "TransformerEncoderBlock")
        _no_attn_dropout(attn_p, training)
        x = _f32c(x)
        bn, t, d = x.shape
        x2d = x.view(bn * t, d)
        x1, s_msa = _msa_fwd(x2d, bn, t, heads, a_ln_w, a_ln_b, w_in, b_in, w_out, b_out, x2d, eps_msa)
        x2, s_mlp, drop = _mlp_fwd(x1, m_ln_w, m_ln_b, w1, b1, w2, b2, float(p) if training else 0.0, x1, eps_mlp)
        ctx.save_for_backward(a_ln_w, m_ln_w, w_in, w_out, w1, w2, *s_msa, *s_mlp)
        ctx.meta = (bn, t, heads, len(s_msa))
        ctx.drop = drop
        return x2.view(bn, t, d)

    @staticmethod
    def backward(ctx, g):
        a_ln_w, m_ln_w, w_in, w_out, w1, w2, *rest = ctx.saved_tensors
        bn, t, heads, n_msa = ctx.meta
        s_msa, s_mlp = tuple(rest[:n_msa]), tuple(rest[n_msa:])
        d = g.shape[-1]
        g2 = _f32c(g).view(bn * t, d)
        g1, (m_dgam, m_dbet, dw1, db1, dw2, db2) = _mlp_bwd(g2, s_mlp, ctx.drop, m_ln_w, w1, w2, has_residual=True)
        g0, (a_dgam, a_dbet, dw_in, db_in, dw_out, db_out) = _msa_bwd(g1, s_msa, bn, t, heads, a_ln_w, w_in, w_out,
                                                                      has_residual=True)
        return (g0.view(bn, t, d), a_dgam, a_dbet, dw_in, db_in, dw_out, db_out, m_dgam, m_dbet, dw1, db1, dw2, db2,
                None, None, None, None, None, None)
