"""Typed Python wrappers over the C ABI: torch tensors in, raw device pointers out.

torch is used here only as the owner of device memory and streams; every function forwards to one
`extern "C"` entry of libb200vit.so and performs no arithmetic itself.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import _C
from ._C import (EPI_ACCUM, EPI_BF16, EPI_BIAS_GELU, EPI_BIAS_RES, EPI_DGELU, EPI_F32, EPI_PATCH, OP_NN, OP_NT,
                 OP_TN)

__all__ = [
    "OP_NT", "OP_NN", "OP_TN", "EPI_BF16", "EPI_BIAS_GELU", "EPI_BIAS_RES", "EPI_DGELU", "EPI_ACCUM", "EPI_F32",
    "EPI_PATCH", "gemm", "cast_bf16", "im2col_patches", "class_token_rows", "embed_param_grads", "patch_grad_rows",
    "layernorm_fwd", "layernorm_bwd", "colsum_bf16", "dropout_cast", "attention_fwd", "attention_bwd", "head_fwd",
    "head_bwd", "softmax_ce", "scale_by_scalar", "l2norm_sq", "l2norm_sq_noalloc", "adam_step", "stream_ptr",
]


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def stream_ptr(device=None):
    """cudaStream_t of torch's current stream on `device` (0 / NULL for the CPU test double). Also tells _C.call which
    device the call belongs to, so that it can guard the launch when that is not the current device."""
    if torch.cuda.is_available():
        idx = getattr(device, "index", None)
        if idx is None:
            idx = torch.cuda.current_device()
        _C.want_device = idx
        return torch.cuda.current_stream(idx).cuda_stream
    return None


def _chk(t: torch.Tensor, dtype, name: str):
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name}: tensor must be contiguous")


_SM_COUNT = {}


def sm_count(device) -> int:
    if not torch.cuda.is_available():
        return 148
    idx = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    if idx not in _SM_COUNT:
        _SM_COUNT[idx] = torch.cuda.get_device_properties(idx).multi_processor_count
    return _SM_COUNT[idx]


def gemm(op: int, epi: int, M: int, N: int, K: int, A: torch.Tensor, lda: int, B: torch.Tensor, ldb: int,
         C: Optional[torch.Tensor], ldc: int, *, C2: Optional[torch.Tensor] = None, ldc2: int = 0,
         bias: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None, ldaux: int = 0,
         dropout_p: float = 0.0, seed: int = 0, offset: int = 0, row_period: int = 0, split_k: int = 1) -> None:
    """split_k: 1 = none; 0 = let the library pick (weight gradients: enough work items for about two waves)."""
    _C.call("b200vit_gemm", op, epi, M, N, K, _ptr(A), lda, _ptr(B), ldb, _ptr(C), ldc, _ptr(C2), ldc2, _ptr(bias),
            _ptr(aux), ldaux, float(dropout_p), int(seed), int(offset), row_period, split_k, stream_ptr(A.device))


def gemm_ln(epi: int, M: int, N: int, K: int, A, lda, B, ldb, C, ldc, *, C2=None, ldc2=0, bias=None, aux=None, ldaux=0,
            dropout_p=0.0, seed=0, offset=0, row_period=0, x16=None, ldx16=0, rowstats=None, ln_stats=None, ln_s=None,
            ln_dim=0, ln_eps=0.0) -> None:
    """b200vit_gemm_ln: forward (NT) GEMM with a LayerNorm-fold epilogue (_C.EPI_RES_STATS .. EPI_LN_GELU)."""
    _C.call("b200vit_gemm_ln", OP_NT, epi, M, N, K, _ptr(A), lda, _ptr(B), ldb, _ptr(C), ldc, _ptr(C2), ldc2, _ptr(bias),
            _ptr(aux), ldaux, float(dropout_p), int(seed), int(offset), row_period, _ptr(x16), ldx16, _ptr(rowstats),
            _ptr(ln_stats), _ptr(ln_s), ln_dim, float(ln_eps), stream_ptr(A.device))


def fold_ln_weight(w, gamma, beta, bias, w16, s, c) -> None:
    rows = w.shape[0]
    cols = w.numel() // rows
    _C.call("b200vit_fold_ln_weight", _ptr(w), _ptr(gamma), _ptr(beta), _ptr(bias), _ptr(w16), _ptr(s), _ptr(c), rows, cols,
            stream_ptr(w.device))


def class_token_rows_ln(cls, pos, x, x16, rowstats, batch, tokens, dim, p, seed, offset) -> None:
    _C.call("b200vit_class_token_rows_ln", _ptr(cls), _ptr(pos), _ptr(x), _ptr(x16), _ptr(rowstats), batch, tokens, dim,
            float(p), int(seed), int(offset), stream_ptr(x.device))


def cast_bf16(src: torch.Tensor, dst: Optional[torch.Tensor] = None) -> torch.Tensor:
    _chk(src, torch.float32, "cast_bf16 src")
    if dst is None:
        dst = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
    _C.call("b200vit_cast_f32_to_bf16", _ptr(src), _ptr(dst), src.numel(), stream_ptr(src.device))
    return dst


def im2col_patches(image: torch.Tensor, patch: int) -> torch.Tensor:
    """image: float32 [B, C, H, W], or uint8 [B, C, H, W] (scaled by 1/255 on the fly)."""
    u8 = image.dtype == torch.uint8
    _chk(image, torch.uint8 if u8 else torch.float32, "image")
    b, c, h, w = image.shape
    out = torch.empty((b * (h // patch) * (w // patch), c * patch * patch), dtype=torch.bfloat16, device=image.device)
    _C.call("b200vit_im2col_patches_u8" if u8 else "b200vit_im2col_patches", _ptr(image), _ptr(out), b, c, h, w, patch,
            stream_ptr(image.device))
    return out


def class_token_rows(cls: torch.Tensor, pos: torch.Tensor, x: torch.Tensor, batch: int, tokens: int, dim: int,
                     p: float, seed: int, offset: int) -> None:
    _C.call("b200vit_class_token_rows", _ptr(cls), _ptr(pos), _ptr(x), batch, tokens, dim, float(p), int(seed),
            int(offset), stream_ptr(x.device))


def embed_param_grads(gx, g_cls, g_pos, g_bias, batch, tokens, dim, p, seed, offset) -> None:
    _C.call("b200vit_embed_param_grads", _ptr(gx), _ptr(g_cls), _ptr(g_pos), _ptr(g_bias), batch, tokens, dim,
            float(p), int(seed), int(offset), stream_ptr(gx.device))


def patch_grad_rows(gx, gy, batch, tokens, dim, p, seed, offset) -> None:
    _C.call("b200vit_patch_grad_rows", _ptr(gx), _ptr(gy), batch, tokens, dim, float(p), int(seed), int(offset),
            stream_ptr(gx.device))


def layernorm_fwd(x, row_stride, gamma, beta, y_bf16, y_f32, mean, rstd, rows, dim, eps) -> None:
    _C.call("b200vit_layernorm_fwd", _ptr(x), row_stride, _ptr(gamma), _ptr(beta), _ptr(y_bf16), _ptr(y_f32),
            _ptr(mean), _ptr(rstd), rows, dim, float(eps), stream_ptr(x.device))


def layernorm_bwd(dy_bf16, dy_f32, x, row_stride, mean, rstd, gamma, gin, gout, gout_stride, gout_bf16, dgamma,
                  dbeta, dbias_next, rows, dim, p=0.0, seed=0, offset=0) -> None:
    _C.call("b200vit_layernorm_bwd", _ptr(dy_bf16), _ptr(dy_f32), _ptr(x), row_stride, _ptr(mean), _ptr(rstd),
            _ptr(gamma), _ptr(gin), _ptr(gout), gout_stride, _ptr(gout_bf16), _ptr(dgamma), _ptr(dbeta),
            _ptr(dbias_next), rows, dim, float(p), int(seed), int(offset), stream_ptr(x.device))


def colsum_bf16(x, ld, out, rows, cols) -> None:
    _C.call("b200vit_colsum_bf16", _ptr(x), ld, _ptr(out), rows, cols, stream_ptr(x.device))


def dropout_cast(g, out_bf16, colsum, rows, cols, p, seed, offset) -> None:
    _C.call("b200vit_dropout_cast", _ptr(g), _ptr(out_bf16), _ptr(colsum), rows, cols, float(p), int(seed),
            int(offset), stream_ptr(g.device))


def attention_fwd(qkv, out, lse, batch, tokens, heads, head_dim, scale, dropout_p=0.0, seed=0, offset=0) -> None:
    """dropout_p > 0: dropout on the attention probabilities (mask = function of (seed, offset))."""
    if dropout_p > 0:
        _C.call("b200vit_attention_fwd_dropout", _ptr(qkv), _ptr(out), _ptr(lse), batch, tokens, heads, head_dim,
                float(scale), float(dropout_p), int(seed), int(offset), stream_ptr(qkv.device))
    else:
        _C.call("b200vit_attention_fwd", _ptr(qkv), _ptr(out), _ptr(lse), batch, tokens, heads, head_dim, float(scale),
                stream_ptr(qkv.device))


def attention_bwd(qkv, out, dout, lse, delta, dqkv, batch, tokens, heads, head_dim, scale, dbias=None, dropout_p=0.0,
                  seed=0, offset=0) -> None:
    """dbias (optional, f32 [3 * heads * head_dim]) receives the column sums of dqkv (the in_proj_bias gradient).
    dropout_p / seed / offset: the triple given to the forward call."""
    if dropout_p > 0:
        _C.call("b200vit_attention_bwd_dropout", _ptr(qkv), _ptr(out), _ptr(dout), _ptr(lse), _ptr(delta), _ptr(dqkv),
                _ptr(dbias), batch, tokens, heads, head_dim, float(scale), float(dropout_p), int(seed), int(offset),
                stream_ptr(qkv.device))
    elif dbias is None:
        _C.call("b200vit_attention_bwd", _ptr(qkv), _ptr(out), _ptr(dout), _ptr(lse), _ptr(delta), _ptr(dqkv), batch,
                tokens, heads, head_dim, float(scale), stream_ptr(qkv.device))
    else:
        _C.call("b200vit_attention_bwd_bias", _ptr(qkv), _ptr(out), _ptr(dout), _ptr(lse), _ptr(delta), _ptr(dqkv),
                _ptr(dbias), batch, tokens, heads, head_dim, float(scale), stream_ptr(qkv.device))


def head_fwd(xn, w, b, logits, batch, dim, classes) -> None:
    _C.call("b200vit_head_fwd", _ptr(xn), _ptr(w), _ptr(b), _ptr(logits), batch, dim, classes, stream_ptr(xn.device))


def head_bwd(xn, w, dlogits, dxn, dw, db, batch, dim, classes) -> None:
    _C.call("b200vit_head_bwd", _ptr(xn), _ptr(w), _ptr(dlogits), _ptr(dxn), _ptr(dw), _ptr(db), batch, dim, classes,
            stream_ptr(xn.device))


def softmax_ce(logits, labels, loss, dlogits, correct, batch, classes) -> None:
    _C.call("b200vit_softmax_ce", _ptr(logits), _ptr(labels), _ptr(loss), _ptr(dlogits), _ptr(correct), batch, classes,
            stream_ptr(logits.device))


def softmax_top1(logits, probs, labels, batch, classes) -> None:
    _C.call("b200vit_softmax_top1", _ptr(logits), _ptr(probs), _ptr(labels), batch, classes, stream_ptr(logits.device))


def scale_by_scalar(src, scalar_dev, dst) -> None:
    _C.call("b200vit_scale_by_scalar", _ptr(src), _ptr(scalar_dev), _ptr(dst), src.numel(), stream_ptr(src.device))


def l2norm_sq(ptrs_dev, sizes_dev, num_tensors, max_size, partials, out_sumsq) -> None:
    _C.call("b200vit_l2norm_sq", _ptr(ptrs_dev), _ptr(sizes_dev), num_tensors, max_size, _ptr(partials),
            partials.numel(), _ptr(out_sumsq), stream_ptr(out_sumsq.device))


def l2norm_sq_noalloc(ptrs_dev, sizes_dev, num_tensors, max_size, out_sumsq) -> None:
    """Gradient sum of squares without any scratch argument (built-in 1 MiB scratch; see include/b200vit.h)."""
    _C.call("b200vit_l2norm_sq_noalloc", _ptr(ptrs_dev), _ptr(sizes_dev), num_tensors, max_size, _ptr(out_sumsq),
            stream_ptr(out_sumsq.device))


# ---- strict-precision ("fp32 mode") entry points ----------------------------------------------------------
def split_bf16x3(src: torch.Tensor, ld_src: int, rows: int, cols: int, dst: torch.Tensor, ld_dst: int,
                 part_stride: int, pattern: int, p: float = 0.0, seed: int = 0, offset: int = 0,
                 row_period: int = 0) -> None:
    _C.call("b200vit_split_bf16x3", _ptr(src), ld_src, rows, cols, _ptr(dst), ld_dst, part_stride, pattern, float(p),
            int(seed), int(offset), row_period, stream_ptr(src.device))


def im2col_patches_f32(image: torch.Tensor, patch: int) -> torch.Tensor:
    _chk(image, torch.float32, "image")
    b, c, h, w = image.shape
    out = torch.empty((b * (h // patch) * (w // patch), c * patch * patch), dtype=torch.float32, device=image.device)
    _C.call("b200vit_im2col_patches_f32", _ptr(image), _ptr(out), b, c, h, w, patch, stream_ptr(image.device))
    return out


def gelu_f32_fwd(a, h, p=0.0, seed=0, offset=0) -> None:
    _C.call("b200vit_gelu_f32_fwd", _ptr(a), _ptr(h), a.numel(), float(p), int(seed), int(offset),
            stream_ptr(a.device))


def gelu_f32_bwd(a, dh, da, p=0.0, seed=0, offset=0) -> None:
    _C.call("b200vit_gelu_f32_bwd", _ptr(a), _ptr(dh), _ptr(da), a.numel(), float(p), int(seed), int(offset),
            stream_ptr(a.device))


def colsum_f32(x, ld, rows, cols, out, p=0.0, seed=0, offset=0, row_period=0) -> None:
    _C.call("b200vit_colsum_f32", _ptr(x), ld, rows, cols, _ptr(out), float(p), int(seed), int(offset), row_period,
            stream_ptr(x.device))


def attention_f32_fwd(qkv, out, lse, batch, tokens, heads, head_dim, scale) -> None:
    _C.call("b200vit_attention_f32_fwd", _ptr(qkv), _ptr(out), _ptr(lse), batch, tokens, heads, head_dim, float(scale),
            stream_ptr(qkv.device))


def attention_f32_bwd(qkv, out, dout, lse, delta, dqkv, batch, tokens, heads, head_dim, scale) -> None:
    _C.call("b200vit_attention_f32_bwd", _ptr(qkv), _ptr(out), _ptr(dout), _ptr(lse), _ptr(delta), _ptr(dqkv), batch,
            tokens, heads, head_dim, float(scale), stream_ptr(qkv.device))


def adam_step(table_dev, num_tensors, max_size, lrs, wds, beta1, beta2, eps, step, decoupled, grad_sumsq, max_norm,
              grad_scale) -> None:
    import ctypes
    n = len(lrs)
    lr_arr = (ctypes.c_float * n)(*[float(v) for v in lrs])
    wd_arr = (ctypes.c_float * n)(*[float(v) for v in wds])
    _C.call("b200vit_adam_step", _ptr(table_dev), num_tensors, max_size, ctypes.cast(lr_arr, ctypes.c_void_p),
            ctypes.cast(wd_arr, ctypes.c_void_p), n, float(beta1), float(beta2), float(eps), int(step),
            int(bool(decoupled)), _ptr(grad_sumsq), float(max_norm), float(grad_scale), stream_ptr(table_dev.device))
