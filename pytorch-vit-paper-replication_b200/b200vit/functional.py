"""Autograd glue between the reference-shaped nn.Modules and the CUDA kernels.

Each torch.autograd.Function below is pure orchestration: it allocates output / saved buffers with torch,
then calls the C ABI (b200vit.kernels). Forward and backward arithmetic all happens in libb200vit.so.

Data layout in HBM (R = B*T token rows, D = embedding dim, M = mlp size, H heads of 64):
  residual stream x ............ f32 [R, D]   (module inputs/outputs are the reference's f32 [B, T, D])
  LayerNorm output ............. bf16 [R, D]  (A operand of the following GEMM, kept for its weight gradient)
  packed q/k/v ................. bf16 [R, 3D] = [B, T, 3, H, 64]; attention reads it in place
  attention output ............. bf16 [R, D]  (+ f32 log-sum-exp [B, H, T])
  fc1 mask*gelu'(pre-act), h ... bf16 [R, M] each (the derivative only when a backward pass will follow)
  weights ...................... f32 master nn.Parameters + bf16 shadows (refreshed by FusedAdam or on demand)
"""
from __future__ import annotations

import math
import os
import weakref
from typing import List, Optional, Tuple

import torch

from . import kernels as K

LN_EPS = 1e-5  # nn.LayerNorm default, reference models/vit.py:83,115,216
HEAD_DIM = 64
MAX_LN_DIM = 1024   # csrc/layernorm.cu: the backward kernels keep a row in registers up to this width

# ---------------------------------------------------------------------------------------------------------
# precision mode: "bf16" (default; bf16 GEMM operands, fp32 accumulation -- logits within 2e-2 of the reference) or
# "fp32" (split-operand GEMMs and fp32 activations, b200vit.fp32 -- logits within 1e-3; 3-4x slower, for parity runs)
# ---------------------------------------------------------------------------------------------------------
_PRECISIONS = ("bf16", "fp32")
_precision = os.environ.get("B200VIT_PRECISION", "bf16")
if _precision not in _PRECISIONS:
    raise RuntimeError(f"B200VIT_PRECISION must be one of {_PRECISIONS}, got {_precision!r}")


def get_precision() -> str:
    return _precision


def set_precision(mode: str) -> str:
    """Select the arithmetic of every b200vit module forward that follows; returns the previous mode."""
    global _precision
    if mode not in _PRECISIONS:
        raise ValueError(f"precision must be one of {_PRECISIONS}, got {mode!r}")
    prev, _precision = _precision, mode
    return prev


class precision:
    """Context manager: `with b200vit.precision("fp32"): logits = model(x)`. The backward of a graph recorded inside
    runs in the same mode wherever it is called."""

    def __init__(self, mode: str):
        self.mode = mode

    def __enter__(self):
        self.prev = set_precision(self.mode)
        return self

    def __exit__(self, *a):
        set_precision(self.prev)

# ---------------------------------------------------------------------------------------------------------
# bf16 shadows of fp32 master weights
# ---------------------------------------------------------------------------------------------------------


class _Shadow:
    __slots__ = ("tensor", "version", "ptr", "owner")

    def __init__(self, tensor, version, ptr, owner):
        self.tensor, self.version, self.ptr, self.owner = tensor, version, ptr, owner


# keyed by id(parameter): tensors cannot be WeakKeyDictionary keys (== on tensors is elementwise)
_SHADOWS: "dict[int, _Shadow]" = {}


def _lookup(p: torch.Tensor) -> Optional[_Shadow]:
    e = _SHADOWS.get(id(p))
    if e is not None and e.owner() is not p:   # id reused by another object after the old parameter died
        del _SHADOWS[id(p)]
        return None
    return e


def bf16_shadow(p: torch.Tensor) -> torch.Tensor:
    """bf16 copy of a weight, re-cast (by our cast kernel) only when the master changed (version / storage)."""
    e = _lookup(p)
    ver, ptr = p._version, p.data_ptr()
    if e is not None and e.version == ver and e.ptr == ptr and e.tensor.device == p.device:
        return e.tensor
    src = p.detach()
    if not src.is_contiguous():
        src = src.contiguous()
    dst = e.tensor if (e is not None and e.tensor.shape == p.shape and e.tensor.device == p.device) else None
    if dst is None:
        # a shadow outlives the call that makes it: allocate it as an ordinary tensor even under torch.inference_mode()
        # (an inference tensor could not be saved for backward by a later training step of the same model)
        with torch.inference_mode(False):
            dst = torch.empty(p.shape, dtype=torch.bfloat16, device=p.device)
    dst = K.cast_bf16(src, dst)
    key = id(p)
    _SHADOWS[key] = _Shadow(dst, ver, ptr, weakref.ref(p, lambda _r, k=key: _SHADOWS.pop(k, None)))
    return dst


def shadow_entry(p: torch.Tensor) -> Optional[_Shadow]:
    return _lookup(p)


def refresh_shadows(model: torch.nn.Module) -> None:
    """Re-cast every bf16 weight shadow of `model` now. Shadows follow a parameter's version counter and storage
    address; writes that bypass the counter -- `p.data.add_(...)` in old-style optimizers, EMA weight swaps through
    `.data`, `dist.broadcast(p.data, ...)` -- leave them stale. Call this after such a write (in-place ops on the
    parameter itself under `torch.no_grad()`, `load_state_dict`, `optimizer.step()` need nothing)."""
    for p in model.parameters():
        e = _lookup(p)
        if e is not None:
            e.version = -1
            bf16_shadow(p)


# ---------------------------------------------------------------------------------------------------------
# dropout seeds: a mask is a function of (seed, offset, element index); offsets count dropout sites
# ---------------------------------------------------------------------------------------------------------
# The counter is the Philox offset of torch's own CUDA generator for the device: `torch.manual_seed` resets it,
# `torch.cuda.get_rng_state` / `set_rng_state` (and therefore checkpoints and torch.utils.checkpoint's RNG
# preservation) capture and replay it, and two runs from the same seed in one process draw the same masks.
# `_dropout_stream` is mixed into the seed: b200vit.dp.DataParallel sets it to the rank so that ranks seeded alike
# (the usual `torch.manual_seed(42)` everywhere) still draw independent masks for their shards of the global batch.
_dropout_stream = 0


def set_dropout_stream(stream: int) -> int:
    """Give this process its own family of dropout masks (the data-parallel rank); returns the previous value."""
    global _dropout_stream
    prev, _dropout_stream = _dropout_stream, int(stream)
    return prev


def next_dropout_key(device=None) -> Tuple[int, int]:
    """(seed, offset) for one dropout site invocation; follows torch.manual_seed() and torch's CUDA RNG state."""
    if torch.cuda.is_available():
        dev = torch.device("cuda") if device is None else torch.device(device)
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        gen = torch.cuda.default_generators[idx]
        seed, off = gen.initial_seed(), gen.get_offset()
        gen.set_offset(off + 4)                     # Philox offsets advance in steps of 4, as torch's own draws do
        count = off // 4 + 1
    else:       # CPU test doubles only: draw the counter from torch's CPU generator, which has no offset to advance
        seed = torch.initial_seed()
        count = int(torch.empty((), dtype=torch.int64).random_(0, 1 << 62))
    seed = (seed ^ (_dropout_stream * 0x9E3779B97F4A7C15)) & 0xFFFFFFFFFFFFFFFF
    return seed, count


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


def _require_cuda(t: torch.Tensor, what: str) -> None:
    if K._C._override is None and not t.is_cuda:
        raise RuntimeError(f"{what}: b200vit runs on CUDA (sm_100a) only; got a {t.device} tensor. "
                           f"There is no CPU fallback -- move the model and inputs to 'cuda'.")


def _empty(shape, dtype, like: torch.Tensor) -> torch.Tensor:
    return torch.empty(shape, dtype=dtype, device=like.device)


def _zeros(shape, like: torch.Tensor) -> torch.Tensor:
    return torch.zeros(shape, dtype=torch.float32, device=like.device)


# ---------------------------------------------------------------------------------------------------------
# Where parameter gradients are written. By default a backward node allocates them; a data-parallel wrapper may
# register, per parameter, a slice of its flat all-reduce buffer as the destination (b200vit.dp): the weight-gradient
# GEMMs then accumulate straight into the communication buffer and `p.grad` is born as a view of it -- no gather copy.
# ---------------------------------------------------------------------------------------------------------
_GRAD_DEST: dict = {}      # id(parameter) -> (weakref to it, callable returning a fresh view of its destination)


def register_grad_destination(p: torch.Tensor, make_view) -> None:
    key = id(p)
    _GRAD_DEST[key] = (weakref.ref(p, lambda _r, k=key: _GRAD_DEST.pop(k, None)), make_view)


def unregister_grad_destination(p: torch.Tensor) -> None:
    _GRAD_DEST.pop(id(p), None)


def grad_out(p: Optional[torch.Tensor], shape, like: torch.Tensor, zero: bool, arena=None) -> torch.Tensor:
    """Buffer for the gradient of parameter `p`: its registered destination when it has one and no gradient has been
    accumulated yet this step (`p.grad is None`; otherwise autograd must add, so a temporary is returned), else the
    next piece of `arena` (zero-filled) or a fresh allocation."""
    if p is not None and p.grad is None:
        e = _GRAD_DEST.get(id(p))
        if e is not None and e[0]() is p:
            t = e[1]()
            if tuple(t.shape) == tuple(shape) and t.device == like.device:
                if zero:
                    t.zero_()
                if arena:
                    arena.skip()
                return t
    if arena:
        return arena.take()
    return _zeros(shape, like) if zero else _empty(shape, torch.float32, like)


class _ZeroArena:
    """One zero-filled fp32 allocation carved into the accumulate-into buffers of a backward node (split-K weight
    gradients, atomically accumulated bias gradients): one fill launch instead of one per buffer. Each piece starts
    on a 256-byte boundary."""

    def __init__(self, like: torch.Tensor, shapes):
        self._views = []
        offs, total = [], 0
        for shp in shapes:
            n = 1
            for v in shp:
                n *= int(v)
            offs.append((total, n, shp))
            total += (n + 63) // 64 * 64
        flat = torch.zeros(total, dtype=torch.float32, device=like.device)
        self._views = [flat[o:o + n].view(shp) for o, n, shp in offs]

    def take(self) -> torch.Tensor:
        return self._views.pop(0)

    def skip(self) -> None:
        self._views.pop(0)


# ---------------------------------------------------------------------------------------------------------
# LayerNorm folded into the GEMMs around it -- forward passes that no backward will follow (no_grad / inference_mode).
#   LN(x) W^T + b = rstd (x (gamma o W)^T - mean s) + c,   s[n] = sum_k gamma[k] W[n,k],   c[n] = sum_k beta[k] W[n,k] + b[n]
# The GEMM that produces the residual stream (patch embedding, out-projection, fc2) also writes it as bf16 and adds up
# each row's (sum, sum of squares); the projection that consumes LayerNorm(x) (in_proj, fc1) multiplies that bf16 copy
# with the gamma-folded weight and applies mean / rstd in its epilogue (csrc/gemm_sm100.cu, epilogues 7-10). The 25
# stand-alone LayerNorm launches of a ViT-B/16 forward disappear. The backward pass wants LN(x) itself as the operand of
# the weight gradients, so training forwards keep the LayerNorm kernel. B200VIT_LN_FOLD=0 / set_ln_fold(False) disable it.
# ---------------------------------------------------------------------------------------------------------
_LN_FOLD = os.environ.get("B200VIT_LN_FOLD", "1") != "0"
# Below this many token rows a forward is bound by the host (one C call per kernel on the folded path, one per block on
# the composite path), not by the 25 LayerNorm launches the fold removes: measured 2.15 vs 1.93 ms at batch 32.
LN_FOLD_MIN_ROWS = int(os.environ.get("B200VIT_LN_FOLD_MIN_ROWS", "12000"))


def set_ln_fold(flag: bool) -> bool:
    global _LN_FOLD
    prev, _LN_FOLD = _LN_FOLD, bool(flag)
    return prev


_FOLDS: dict = {}      # id(weight) -> [folded bf16 weight, s, c, stamp, weakrefs of (w, gamma, beta, bias)]


def _fold_stamp(w, gamma, beta, bias):
    return (w._version, gamma._version, beta._version, bias._version, w.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
            bias.data_ptr())


def folded_weight(w: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, bias: torch.Tensor):
    """(bf16(gamma o W), s, c) for a Linear fed by LayerNorm(gamma, beta); re-made -- in place, so that a recorded
    CUDA graph keeps seeing it -- when any of the four changed."""
    key = id(w)
    stamp = _fold_stamp(w, gamma, beta, bias)
    e = _FOLDS.get(key)
    if e is not None and e[4][0]() is w and e[0].device == w.device and e[0].shape == w.shape:
        if e[3] != stamp:
            K.fold_ln_weight(w.detach().contiguous(), gamma.detach(), beta.detach(), bias.detach(), e[0], e[1], e[2])
            e[3] = stamp
            e[4] = (e[4][0], weakref.ref(gamma), weakref.ref(beta), weakref.ref(bias))
        return e[0], e[1], e[2]
    n = w.shape[0]
    w16 = torch.empty(w.shape, dtype=torch.bfloat16, device=w.device)
    sv = torch.empty(n, dtype=torch.float32, device=w.device)
    cv = torch.empty(n, dtype=torch.float32, device=w.device)
    K.fold_ln_weight(w.detach().contiguous(), gamma.detach(), beta.detach(), bias.detach(), w16, sv, cv)
    _FOLDS[key] = [w16, sv, cv, stamp, (weakref.ref(w, lambda _r, k=key: _FOLDS.pop(k, None)), weakref.ref(gamma),
                                         weakref.ref(beta), weakref.ref(bias))]
    return w16, sv, cv


def refresh_folds() -> None:
    """Re-make (in place) every folded weight whose sources changed since it was made: what a CUDA-graph replay of a
    folded forward needs after an in-place parameter update (b200vit.inference.GraphedForward calls it)."""
    for e in list(_FOLDS.values()):
        w, gamma, beta, bias = (r() for r in e[4])
        if w is None or gamma is None or beta is None or bias is None:
            continue
        if e[3] != _fold_stamp(w, gamma, beta, bias) and e[0].device == w.device:
            folded_weight(w, gamma, beta, bias)


def _ln_side() -> dict:     # data_ptr of a residual-stream tensor -> (its bf16 copy, row statistics, version, the tensor)
    d = getattr(_TLS, "ln_side", None)
    if d is None:
        d = _TLS.ln_side = {}
    return d


def _version_of(x: torch.Tensor) -> int:
    return -1 if x.is_inference() else x._version      # tensors made under inference_mode carry no version counter


def _ln_side_put(x: torch.Tensor, x16: torch.Tensor, stats: torch.Tensor) -> None:
    d = _ln_side()
    d.clear()                   # one stream in flight per thread: whatever was not consumed is stale
    d[x.data_ptr()] = (x16, stats, _version_of(x), x)


def _ln_side_take(x: torch.Tensor):
    e = _ln_side().pop(x.data_ptr(), None)
    if e is not None and e[2] == _version_of(x) and e[3].shape == x.shape:
        return e[0], e[1]
    return None


def _block_forward_folded(x, x2d, bn, t, heads, params, shadows, p, attn_p, eps_msa, eps_mlp):
    """Encoder block forward with both LayerNorms folded (no activations are kept)."""
    from . import _C
    a_ln_w, a_ln_b, w_in, b_in, w_out, b_out, m_ln_w, m_ln_b, w1, b1, w2, b2 = params
    w_in16, w_out16, w1_16, w2_16 = shadows
    r, d = x2d.shape
    m = w1_16.shape[0]
    if d != heads * HEAD_DIM:
        raise NotImplementedError(f"b200vit attention kernels need head_dim == 64 (embedding_dim {d}, heads {heads})")
    dev = x2d.device
    key1 = next_dropout_key(dev) if p > 0 else (0, 0)
    key2 = next_dropout_key(dev) if p > 0 else (0, 0)
    akey = next_dropout_key(dev) if attn_p > 0 else (0, 0)
    side = _ln_side_take(x)
    qkv = _empty((r, 3 * d), torch.bfloat16, x2d)
    if side is not None:
        x16, stats = side
        wf, sv, cv = folded_weight(w_in, a_ln_w, a_ln_b, b_in)
        K.gemm_ln(_C.EPI_LN_BF16, r, 3 * d, d, x16, d, wf, d, qkv, 3 * d, bias=cv, ln_stats=stats, ln_s=sv, ln_dim=d,
                  ln_eps=eps_msa)
    else:                       # the stream did not come from one of our producers: normalise it the ordinary way
        ln = _empty((r, d), torch.bfloat16, x2d)
        mean = _empty((r,), torch.float32, x2d)
        rstd = _empty((r,), torch.float32, x2d)
        K.layernorm_fwd(x2d, d, a_ln_w, a_ln_b, ln, None, mean, rstd, r, d, eps_msa)
        K.gemm(K.OP_NT, K.EPI_BF16, r, 3 * d, d, ln, d, w_in16, d, qkv, 3 * d, bias=b_in)
    o = _empty((r, d), torch.bfloat16, x2d)
    lse = _empty((bn, heads, t), torch.float32, x2d)
    K.attention_fwd(qkv, o, lse, bn, t, heads, HEAD_DIM, 1.0 / math.sqrt(HEAD_DIM), attn_p, akey[0], akey[1])
    x1 = _empty((r, d), torch.float32, x2d)
    x1_16 = _empty((r, d), torch.bfloat16, x2d)
    st1 = _zeros((r, 2), x2d)
    K.gemm_ln(_C.EPI_RES_STATS, r, d, d, o, d, w_out16, d, x1, d, bias=b_out, aux=x2d, ldaux=d, x16=x1_16, ldx16=d,
              rowstats=st1)
    wf, sv, cv = folded_weight(w1, m_ln_w, m_ln_b, b1)
    h = _empty((r, m), torch.bfloat16, x2d)
    K.gemm_ln(_C.EPI_LN_GELU, r, m, d, x1_16, d, wf, d, None, m, C2=h, ldc2=m, bias=cv, dropout_p=p, seed=key1[0],
              offset=key1[1], ln_stats=st1, ln_s=sv, ln_dim=d, ln_eps=eps_mlp)
    x2 = _empty((r, d), torch.float32, x2d)
    x2_16 = _empty((r, d), torch.bfloat16, x2d)
    st2 = _zeros((r, 2), x2d)
    K.gemm_ln(_C.EPI_RES_STATS, r, d, m, h, m, w2_16, m, x2, d, bias=b2, aux=x1, ldaux=d, dropout_p=p, seed=key2[0],
              offset=key2[1], x16=x2_16, ldx16=d, rowstats=st2)
    out = x2.view(bn, t, d)
    _ln_side_put(out, x2_16, st2)
    return out


# ---------------------------------------------------------------------------------------------------------
# Patch embedding  (reference PatchEmbedding.forward, models/vit.py:45-67)
# ---------------------------------------------------------------------------------------------------------
class PatchEmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, w, b, cls, pos, w16, patch: int, p_drop: float, training: bool, grad_mode: bool = True):
        _require_cuda(img, "PatchEmbedding")
        # uint8 images are accepted as they are (converted to [0, 1] inside the im2col kernel: the reference's
        # ToDtype(float32, scale=True) on the device, a 4x smaller host->device copy); anything else becomes fp32
        img = img.contiguous() if img.dtype == torch.uint8 else _f32c(img)
        bn, c, h, wd = img.shape
        d = w.shape[0]
        np_ = (h // patch) * (wd // patch)
        t = np_ + 1
        if pos.shape[1] != t:
            raise RuntimeError(f"image of {h}x{wd} gives {t} tokens but position_embedding has {pos.shape[1]}")
        kdim = c * patch * patch
        patches = K.im2col_patches(img, patch)
        x = _empty((bn, t, d), torch.float32, img)
        p = float(p_drop) if training else 0.0
        seed, off = next_dropout_key(img.device) if p > 0 else (0, 0)
        if _LN_FOLD and bn * t >= LN_FOLD_MIN_ROWS and not (bool(grad_mode) and any(ctx.needs_input_grad)):
            # no backward will follow: also emit what the first encoder block's folded LayerNorm needs
            from . import _C
            x16 = _empty((bn, t, d), torch.bfloat16, img)
            stats = _zeros((bn * t, 2), img)
            K.gemm_ln(_C.EPI_PATCH_STATS, bn * np_, d, kdim, patches, kdim, w16, kdim, x, d, bias=b, aux=pos, ldaux=d,
                      dropout_p=p, seed=seed, offset=off, row_period=np_, x16=x16, ldx16=d, rowstats=stats)
            K.class_token_rows_ln(cls, pos, x, x16, stats, bn, t, d, p, seed, off)
            _ln_side_put(x, x16, stats)
            ctx.meta = None
            return x
        K.gemm(K.OP_NT, K.EPI_PATCH, bn * np_, d, kdim, patches, kdim, w16, kdim, x, d, bias=b, aux=pos, ldaux=d,
               dropout_p=p, seed=seed, offset=off, row_period=np_)
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
        gy = _empty((rows, d), torch.bfloat16, gx)
        K.patch_grad_rows(gx, gy, bn, t, d, p, seed, off)
        gw = _zeros((d, kdim), gx)
        K.gemm(K.OP_TN, K.EPI_ACCUM, d, kdim, rows, gy, d, patches, kdim, gw, kdim,
               split_k=0)
        return None, gw.view(wshape), g_b, g_cls, g_pos, None, None, None, None, None


# ---------------------------------------------------------------------------------------------------------
# MSA / MLP halves of an encoder block (reference models/vit.py:92-98, 128-131, 166-169)
# ---------------------------------------------------------------------------------------------------------
# ---------------------------------------------------------------------------------------------------------
# Cross-block hand-over in the backward pass. The gradient leaving encoder block i+1 (fp32 residual-stream gradient,
# written by the LayerNorm backward of its attention half) is consumed by block i's MLP backward, which first needs
# bf16(dropout_mask_i * g) and its column sums (fc2 backward operands; the Dropout of models/vit.py:125). The
# LayerNorm backward kernel can emit both in the pass that writes g, sparing a 3-pass elementwise kernel per layer.
# The two blocks are separate autograd nodes, so the hand-over is keyed by the storage address of g:
#   forward : block i records (address of its output, its fc2-dropout p and key); block i+1, whose input is that
#             very tensor, picks the triple up;
#   backward: block i+1 emits the extra outputs with block i's mask and files them under g's address; block i looks
#             its incoming gradient up and falls back to the stand-alone kernel if anything does not match (another
#             consumer, a hook that modified g, a different mask).
# ---------------------------------------------------------------------------------------------------------
# Both records are PER THREAD (a forward runs on the caller's thread, a backward on one autograd thread: producer and
# consumer of a record always share a thread), so models trained from different threads cannot see each other's. A side
# entry keeps a strong reference to the gradient tensor it belongs to: while the entry exists the allocator cannot hand
# that address to another tensor, so a stale entry can never be mistaken for a fresh one; entries left over by a backward
# that was not run to the end are dropped at the next forward.
import threading

_TLS = threading.local()


def _grad_side() -> dict:       # data_ptr of g -> (g16, colsum, p, (seed, offset), version of g, g itself)
    d = getattr(_TLS, "grad_side", None)
    if d is None:
        d = _TLS.grad_side = {}
    return d


def _side_put(gx: torch.Tensor, gx16, dbn, p: float, key) -> None:
    _grad_side()[gx.data_ptr()] = (gx16, dbn, p, key, gx._version, gx)


def _side_take(g: torch.Tensor, p: float, key, shape):
    side = _grad_side().pop(g.data_ptr(), None)
    if side is not None and side[2] == p and side[3] == tuple(key) and side[4] == g._version and \
            tuple(side[0].shape) == tuple(shape):
        return side
    return None


def _note_block_output(y: torch.Tensor, p: float, key) -> None:
    _TLS.last_block_out = (y.data_ptr(), float(p), tuple(key))
    side = getattr(_TLS, "grad_side", None)
    if side:                    # left over from a backward pass that never reached its consumer
        side.clear()


def _producer_dropout(x: torch.Tensor):
    """(p, key) of the encoder block that produced x, if x is that block's output tensor; else None."""
    rec = getattr(_TLS, "last_block_out", None)
    _TLS.last_block_out = None      # one consumer per record; a stale address must not match later
    if rec is not None and rec[0] == x.data_ptr():
        return rec[1], rec[2]
    return None


def _msa_forward(x2d, bn, t, heads, ln_w, ln_b, w_in16, b_in, w_out16, b_out, residual, attn_p=0.0, eps=LN_EPS):
    r, d = x2d.shape
    if d != heads * HEAD_DIM:
        raise NotImplementedError(f"b200vit attention kernels need head_dim == 64 (embedding_dim {d}, heads {heads})")
    ln = _empty((r, d), torch.bfloat16, x2d)
    mean = _empty((r,), torch.float32, x2d)
    rstd = _empty((r,), torch.float32, x2d)
    K.layernorm_fwd(x2d, d, ln_w, ln_b, ln, None, mean, rstd, r, d, eps)
    qkv = _empty((r, 3 * d), torch.bfloat16, x2d)
    K.gemm(K.OP_NT, K.EPI_BF16, r, 3 * d, d, ln, d, w_in16, d, qkv, 3 * d, bias=b_in)
    o = _empty((r, d), torch.bfloat16, x2d)
    lse = _empty((bn, heads, t), torch.float32, x2d)
    akey = next_dropout_key(x2d.device) if attn_p > 0 else (0, 0)      # dropout on the attention probabilities (vit.py:86-90)
    K.attention_fwd(qkv, o, lse, bn, t, heads, HEAD_DIM, 1.0 / math.sqrt(HEAD_DIM), attn_p, akey[0], akey[1])
    y = _empty((r, d), torch.float32, x2d)
    K.gemm(K.OP_NT, K.EPI_BIAS_RES, r, d, d, o, d, w_out16, d, y, d, bias=b_out, aux=residual, ldaux=d)
    return y, (x2d, mean, rstd, ln, qkv, o, lse), (attn_p, akey)


def _msa_backward(g, saved, bn, t, heads, ln_w, w_in16, w_out16, has_residual, g16=None, db_out=None,
                  adrop=(0.0, (0, 0)), arena=None, consumer_drop=None, params=None):
    # params (optional): the Parameters (ln_w, ln_b, w_in, b_in, w_out, b_out) whose gradients are produced here; when a
    # gradient destination is registered for one of them (data parallel) its gradient is written there directly
    x2d, mean, rstd, ln, qkv, o, lse = saved
    r, d = x2d.shape
    dev = x2d.device
    P = params if params is not None else (None,) * 6
    if g16 is None:
        g16 = _empty((r, d), torch.bfloat16, x2d)
        db_out = grad_out(P[5], (d,), x2d, zero=False)
        K.dropout_cast(g, g16, db_out, r, d, 0.0, 0, 0)
    if arena is None:
        arena = _ZeroArena(x2d, [(d, d), (3 * d, d)])
    dw_out = grad_out(P[4], (d, d), x2d, zero=True, arena=arena)
    K.gemm(K.OP_TN, K.EPI_ACCUM, d, d, r, g16, d, o, d, dw_out, d, split_k=0)
    do = _empty((r, d), torch.bfloat16, x2d)
    K.gemm(K.OP_NN, K.EPI_BF16, r, d, d, g16, d, w_out16, d, do, d)
    delta = _empty((bn, heads, t), torch.float32, x2d)
    dqkv = _empty((r, 3 * d), torch.bfloat16, x2d)
    db_in = grad_out(P[3], (3 * d,), x2d, zero=False)  # in_proj_bias gradient: column sums of dqkv, from the same kernel
    K.attention_bwd(qkv, o, do, lse, delta, dqkv, bn, t, heads, HEAD_DIM, 1.0 / math.sqrt(HEAD_DIM), dbias=db_in,
                    dropout_p=adrop[0], seed=adrop[1][0], offset=adrop[1][1])
    dw_in = grad_out(P[2], (3 * d, d), x2d, zero=True, arena=arena)
    K.gemm(K.OP_TN, K.EPI_ACCUM, 3 * d, d, r, dqkv, 3 * d, ln, d, dw_in, d, split_k=0)
    dln = _empty((r, d), torch.bfloat16, x2d)
    K.gemm(K.OP_NN, K.EPI_BF16, r, d, 3 * d, dqkv, 3 * d, w_in16, d, dln, d)
    gx = _empty((r, d), torch.float32, x2d)
    dgamma = grad_out(P[0], (d,), x2d, zero=False)
    dbeta = grad_out(P[1], (d,), x2d, zero=False)
    if consumer_drop is None:
        K.layernorm_bwd(dln, None, x2d, d, mean, rstd, ln_w, g if has_residual else None, gx, d, None, dgamma, dbeta,
                        None, r, d)
    else:
        # also emit what the previous block's MLP backward needs first: bf16(mask * gx) with ITS fc2-dropout mask
        cp, ckey = consumer_drop
        gx16 = _empty((r, d), torch.bfloat16, x2d)
        dbn = _empty((d,), torch.float32, x2d)
        K.layernorm_bwd(dln, None, x2d, d, mean, rstd, ln_w, g if has_residual else None, gx, d, gx16, dgamma, dbeta,
                        dbn, r, d, cp, ckey[0], ckey[1])
        _side_put(gx, gx16, dbn, cp, ckey)
    return gx, (dgamma, dbeta, dw_in, db_in, dw_out, db_out)


def _mlp_forward(x2d, ln_w, ln_b, w1_16, b1, w2_16, b2, p, residual, keep_preact, eps=LN_EPS):
    r, d = x2d.shape
    m = w1_16.shape[0]
    ln = _empty((r, d), torch.bfloat16, x2d)
    mean = _empty((r,), torch.float32, x2d)
    rstd = _empty((r,), torch.float32, x2d)
    K.layernorm_fwd(x2d, d, ln_w, ln_b, ln, None, mean, rstd, r, d, eps)
    key1 = next_dropout_key(x2d.device) if p > 0 else (0, 0)
    key2 = next_dropout_key(x2d.device) if p > 0 else (0, 0)
    a = _empty((r, m), torch.bfloat16, x2d) if keep_preact else None
    h = _empty((r, m), torch.bfloat16, x2d)
    K.gemm(K.OP_NT, K.EPI_BIAS_GELU, r, m, d, ln, d, w1_16, d, a, m, C2=h, ldc2=m, bias=b1, dropout_p=p,
           seed=key1[0], offset=key1[1])
    y = _empty((r, d), torch.float32, x2d)
    K.gemm(K.OP_NT, K.EPI_BIAS_RES, r, d, m, h, m, w2_16, m, y, d, bias=b2, aux=residual, ldaux=d, dropout_p=p,
           seed=key2[0], offset=key2[1])
    return y, (x2d, mean, rstd, ln, a, h), (p, key1, key2)


def _mlp_backward(g, saved, drop, ln_w, w1_16, w2_16, has_residual, want_bf16_out, arena=None, params=None,
                  next_bias_param=None):
    # params (optional): the Parameters (ln_w, ln_b, w1, b1, w2, b2); next_bias_param: the bias whose gradient the
    # LayerNorm backward below emits on the side (out_proj.bias of the same block) -- see _msa_backward
    x2d, mean, rstd, ln, a, h = saved
    p, key1, key2 = drop
    r, d = x2d.shape
    m = w1_16.shape[0]
    dev = x2d.device
    P = params if params is not None else (None,) * 6
    side = _side_take(g, p, key2, (r, d))
    if side is not None:
        g16, db2 = side[0], side[1]     # produced by the next block's LayerNorm backward together with g
    else:
        g16 = _empty((r, d), torch.bfloat16, x2d)
        db2 = grad_out(P[5], (d,), x2d, zero=False)
        K.dropout_cast(g, g16, db2, r, d, p, key2[0], key2[1])
    if arena is None:
        arena = _ZeroArena(x2d, [(d, m), (m,), (m, d)])
    dw2 = grad_out(P[4], (d, m), x2d, zero=True, arena=arena)
    K.gemm(K.OP_TN, K.EPI_ACCUM, d, m, r, g16, d, h, m, dw2, m, split_k=0)
    da = _empty((r, m), torch.bfloat16, x2d)
    db1 = grad_out(P[3], (m,), x2d, zero=True, arena=arena)
    # a = mask * gelu'(pre-activation); the epilogue also accumulates the fc1 bias gradient (column sums of da) into db1
    K.gemm(K.OP_NN, K.EPI_DGELU, r, m, d, g16, d, w2_16, m, da, m, bias=db1, aux=a, ldaux=m)
    dw1 = grad_out(P[2], (m, d), x2d, zero=True, arena=arena)
    K.gemm(K.OP_TN, K.EPI_ACCUM, m, d, r, da, m, ln, d, dw1, d, split_k=0)
    dln = g16  # reuse: g16 is dead after the two GEMMs above (same shape / dtype)
    K.gemm(K.OP_NN, K.EPI_BF16, r, d, m, da, m, w1_16, d, dln, d)
    gx = _empty((r, d), torch.float32, x2d)
    dgamma = grad_out(P[0], (d,), x2d, zero=False)
    dbeta = grad_out(P[1], (d,), x2d, zero=False)
    gx16 = _empty((r, d), torch.bfloat16, x2d) if want_bf16_out else None
    dbn = grad_out(next_bias_param, (d,), x2d, zero=False) if want_bf16_out else None
    K.layernorm_bwd(dln, None, x2d, d, mean, rstd, ln_w, g if has_residual else None, gx, d, gx16, dgamma, dbeta,
                    dbn, r, d)
    return gx, (dgamma, dbeta, dw1, db1, dw2, db2), gx16, dbn


class MSAFn(torch.autograd.Function):
    """MultiHeadSelfAttentionBlock.forward (models/vit.py:92-98): LN -> self-attention, no residual."""

    @staticmethod
    def forward(ctx, x, ln_w, ln_b, w_in, b_in, w_out, b_out, w_in16, w_out16, heads: int, attn_p: float = 0.0,
                training: bool = False, eps: float = LN_EPS):
        _require_cuda(x, "MultiHeadSelfAttentionBlock")
        x = _f32c(x)
        bn, t, d = x.shape
        y, saved, adrop = _msa_forward(x.view(bn * t, d), bn, t, heads, ln_w, ln_b, w_in16, b_in, w_out16, b_out, None,
                                       float(attn_p) if training else 0.0, eps)
        ctx.save_for_backward(ln_w, w_in16, w_out16, *saved)
        ctx.meta = (bn, t, heads)
        ctx.adrop = adrop
        return y.view(bn, t, d)

    @staticmethod
    def backward(ctx, g):
        ln_w, w_in16, w_out16, *saved = ctx.saved_tensors
        bn, t, heads = ctx.meta
        d = g.shape[-1]
        gx, (dgam, dbet, dw_in, db_in, dw_out, db_out) = _msa_backward(
            _f32c(g).view(bn * t, d), tuple(saved), bn, t, heads, ln_w, w_in16, w_out16, has_residual=False,
            adrop=ctx.adrop)
        return gx.view(bn, t, d), dgam, dbet, dw_in, db_in, dw_out, db_out, None, None, None, None, None, None


class MLPFn(torch.autograd.Function):
    """MLPBlock.forward (models/vit.py:128-131): LN -> Linear -> GELU -> Dropout -> Linear -> Dropout."""

    @staticmethod
    def forward(ctx, x, ln_w, ln_b, w1, b1, w2, b2, w1_16, w2_16, p: float, training: bool, grad_mode: bool = True,
                eps: float = LN_EPS):
        # grad_mode: torch.is_grad_enabled() as seen by the CALLER (inside forward it is always False, and
        # needs_input_grad stays True under no_grad / inference_mode): without it, inference would save activations
        _require_cuda(x, "MLPBlock")
        x = _f32c(x)
        shape = x.shape
        d = shape[-1]
        x2d = x.view(-1, d)
        needs_bwd = bool(grad_mode) and any(ctx.needs_input_grad)
        y, saved, drop = _mlp_forward(x2d, ln_w, ln_b, w1_16, b1, w2_16, b2, float(p) if training else 0.0, None,
                                      keep_preact=needs_bwd, eps=eps)
        if needs_bwd:
            ctx.save_for_backward(ln_w, w1_16, w2_16, *saved)
        ctx.drop = drop
        return y.view(shape)

    @staticmethod
    def backward(ctx, g):
        ln_w, w1_16, w2_16, *saved = ctx.saved_tensors
        shape = g.shape
        d = shape[-1]
        gx, (dgam, dbet, dw1, db1, dw2, db2), _, _ = _mlp_backward(
            _f32c(g).view(-1, d), tuple(saved), ctx.drop, ln_w, w1_16, w2_16, has_residual=False, want_bf16_out=False)
        return gx.view(shape), dgam, dbet, dw1, db1, dw2, db2, None, None, None, None, None, None


# ---------------------------------------------------------------------------------------------------------
# Encoder block through the composite C entry points: ONE Python -> C transition per block per direction
# (b200vit_encoder_block_fwd / _bwd, csrc/block.cu) instead of one per kernel, and two allocations per direction
# instead of one per buffer. B200VIT_COMPOSITE=0 falls back to driving the kernels one by one (the helpers above).
# ---------------------------------------------------------------------------------------------------------
_COMPOSITE = os.environ.get("B200VIT_COMPOSITE", "1") != "0"


def set_composite(flag: bool) -> bool:
    """Drive encoder blocks through the composite C calls (True, default) or kernel by kernel (False: what the
    per-kernel timers of bench.py need); returns the previous setting."""
    global _COMPOSITE
    prev, _COMPOSITE = _COMPOSITE, bool(flag)
    return prev


def _carve(like: torch.Tensor, dtype, sizes):
    """One allocation cut into 1-D pieces of `sizes` elements, each starting on a 256-byte boundary."""
    offs, total = [], 0
    for n in sizes:
        offs.append(total)
        total += (int(n) + 127) // 128 * 128
    flat = torch.empty(total, dtype=dtype, device=like.device)
    return [flat[o:o + int(n)] for o, n in zip(offs, sizes)]


def _block_forward_c(x2d, bn, t, heads, params, shadows, p, attn_p, eps_msa, eps_mlp, needs_bwd):
    from . import _C
    a_ln_w, a_ln_b, w_in, b_in, w_out, b_out, m_ln_w, m_ln_b, w1, b1, w2, b2 = params
    w_in16, w_out16, w1_16, w2_16 = shadows
    r, d = x2d.shape
    m = w1_16.shape[0]
    if d != heads * HEAD_DIM:
        raise NotImplementedError(f"b200vit attention kernels need head_dim == 64 (embedding_dim {d}, heads {heads})")
    dev = x2d.device
    key1 = next_dropout_key(dev) if p > 0 else (0, 0)
    key2 = next_dropout_key(dev) if p > 0 else (0, 0)
    akey = next_dropout_key(dev) if attn_p > 0 else (0, 0)
    ln1, qkv, o, ln2, a, h = _carve(x2d, torch.bfloat16, [r * d, 3 * r * d, r * d, r * d, r * m if needs_bwd else 0, r * m])
    mean1, rstd1, lse, x1, mean2, rstd2 = _carve(x2d, torch.float32, [r, r, bn * heads * t, r * d, r, r])
    x2 = torch.empty((r, d), dtype=torch.float32, device=dev)
    A = _C.BlockFwdArgs()
    A.struct_bytes = _C.C.sizeof(_C.BlockFwdArgs)
    A.batch, A.tokens, A.heads, A.dim, A.mlp = bn, t, heads, d, m
    A.eps_msa, A.eps_mlp, A.mlp_p, A.attn_p = eps_msa, eps_mlp, p, attn_p
    A.key1_seed, A.key1_offset, A.key2_seed, A.key2_offset = key1[0], key1[1], key2[0], key2[1]
    A.attn_seed, A.attn_offset = akey[0], akey[1]
    A.x = x2d.data_ptr()
    A.ln1_w, A.ln1_b, A.b_in, A.b_out = a_ln_w.data_ptr(), a_ln_b.data_ptr(), b_in.data_ptr(), b_out.data_ptr()
    A.ln2_w, A.ln2_b, A.b1, A.b2 = m_ln_w.data_ptr(), m_ln_b.data_ptr(), b1.data_ptr(), b2.data_ptr()
    A.w_in16, A.w_out16, A.w1_16, A.w2_16 = w_in16.data_ptr(), w_out16.data_ptr(), w1_16.data_ptr(), w2_16.data_ptr()
    A.ln1, A.qkv, A.o, A.ln2, A.h = ln1.data_ptr(), qkv.data_ptr(), o.data_ptr(), ln2.data_ptr(), h.data_ptr()
    A.a = a.data_ptr() if needs_bwd else None
    A.mean1, A.rstd1, A.lse, A.x1 = mean1.data_ptr(), rstd1.data_ptr(), lse.data_ptr(), x1.data_ptr()
    A.mean2, A.rstd2, A.x2 = mean2.data_ptr(), rstd2.data_ptr(), x2.data_ptr()
    _C.call("b200vit_encoder_block_fwd", _C.C.byref(A), K.stream_ptr(dev))
    saved = (x2d, mean1, rstd1, ln1, qkv, o, lse, x1, mean2, rstd2, ln2, a, h)
    return x2, saved, (p, key1, key2), (attn_p, akey)


def _block_backward_c(g2, saved, bn, t, heads, weights, drop, adrop, consumer_drop, grad_params):
    from . import _C
    x2d, mean1, rstd1, ln1, qkv, o, lse, x1, mean2, rstd2, ln2, a, h = saved
    a_ln_w, m_ln_w, w_in16, w_out16, w1_16, w2_16 = weights
    p, key1, key2 = drop
    attn_p, akey = adrop
    r, d = x2d.shape
    m = w1_16.shape[0]
    dev = x2d.device
    side = _side_take(g2, p, key2, (r, d))
    handed = side is not None
    g16, da, dout, dqkv, g1_16 = _carve(x2d, torch.bfloat16, [0 if handed else r * d, r * m, r * d, 3 * r * d, r * d])
    g1, delta = _carve(x2d, torch.float32, [r * d, bn * heads * t])
    gx = torch.empty((r, d), dtype=torch.float32, device=dev)
    gp = grad_params
    shapes = [(d,), (d,), (3 * d, d), (3 * d,), (d, d), (d,), (d,), (d,), (m, d), (m,), (d, m), (d,)]
    if any(id(q) in _GRAD_DEST for q in gp):
        grads = [grad_out(q, shp, x2d, zero=False) for q, shp in zip(gp, shapes)]
    else:
        flat = _carve(x2d, torch.float32, [math.prod(shp) for shp in shapes])
        grads = [f.view(shp) for f, shp in zip(flat, shapes)]
    if handed:
        g16, grads[11] = side[0], side[1]       # bf16(mask2 * g) and the fc2 bias gradient came with g
    gx16 = dbn = None
    if consumer_drop is not None:
        gx16 = torch.empty((r, d), dtype=torch.bfloat16, device=dev)
        dbn = torch.empty((d,), dtype=torch.float32, device=dev)
    A = _C.BlockBwdArgs()
    A.struct_bytes = _C.C.sizeof(_C.BlockBwdArgs)
    A.batch, A.tokens, A.heads, A.dim, A.mlp = bn, t, heads, d, m
    A.mlp_p, A.attn_p = p, attn_p
    A.key1_seed, A.key1_offset, A.key2_seed, A.key2_offset = key1[0], key1[1], key2[0], key2[1]
    A.attn_seed, A.attn_offset = akey[0], akey[1]
    A.g, A.g16_ready, A.zero_accumulators = g2.data_ptr(), int(handed), 1
    A.x, A.mean1, A.rstd1, A.lse, A.x1 = x2d.data_ptr(), mean1.data_ptr(), rstd1.data_ptr(), lse.data_ptr(), x1.data_ptr()
    A.mean2, A.rstd2 = mean2.data_ptr(), rstd2.data_ptr()
    A.ln1, A.qkv, A.o, A.ln2, A.a, A.h = ln1.data_ptr(), qkv.data_ptr(), o.data_ptr(), ln2.data_ptr(), a.data_ptr(), h.data_ptr()
    A.ln1_w, A.ln2_w = a_ln_w.data_ptr(), m_ln_w.data_ptr()
    A.w_in16, A.w_out16, A.w1_16, A.w2_16 = w_in16.data_ptr(), w_out16.data_ptr(), w1_16.data_ptr(), w2_16.data_ptr()
    A.g16, A.da, A.dout, A.dqkv, A.g1_16 = g16.data_ptr(), da.data_ptr(), dout.data_ptr(), dqkv.data_ptr(), g1_16.data_ptr()
    A.g1, A.delta, A.gx = g1.data_ptr(), delta.data_ptr(), gx.data_ptr()
    (A.d_ln1_w, A.d_ln1_b, A.dw_in, A.db_in, A.dw_out, A.db_out, A.d_ln2_w, A.d_ln2_b, A.dw1, A.db1, A.dw2,
     A.db2) = [t_.data_ptr() for t_ in grads]
    if consumer_drop is not None:
        cp, ckey = consumer_drop
        A.gx16_out, A.dbn_out, A.consumer_p = gx16.data_ptr(), dbn.data_ptr(), cp
        A.consumer_seed, A.consumer_offset = ckey[0], ckey[1]
    _C.call("b200vit_encoder_block_bwd", _C.C.byref(A), K.stream_ptr(dev))
    if consumer_drop is not None:
        _side_put(gx, gx16, dbn, consumer_drop[0], consumer_drop[1])
    return gx, grads


class EncoderBlockFn(torch.autograd.Function):
    """TransformerEncoderBlock.forward (models/vit.py:166-169) as one autograd node:
    x1 = msa(x) + x ; x2 = mlp(x1) + x1, with both residual adds fused into GEMM epilogues."""

    @staticmethod
    def forward(ctx, x, a_ln_w, a_ln_b, w_in, b_in, w_out, b_out, m_ln_w, m_ln_b, w1, b1, w2, b2,
                w_in16, w_out16, w1_16, w2_16, heads: int, p: float, training: bool, attn_p: float = 0.0,
                grad_mode: bool = True, eps_msa: float = LN_EPS, eps_mlp: float = LN_EPS):
        _require_cuda(x, "TransformerEncoderBlock")
        # the Parameters themselves (not saved tensors): backward asks functional.grad_out where each gradient goes
        ctx.grad_params = (a_ln_w, a_ln_b, w_in, b_in, w_out, b_out, m_ln_w, m_ln_b, w1, b1, w2, b2)
        x = _f32c(x)
        bn, t, d = x.shape
        x2d = x.view(bn * t, d)
        needs_bwd = bool(grad_mode) and any(ctx.needs_input_grad)
        ctx.consumer_drop = _producer_dropout(x) if needs_bwd else None      # see "cross-block hand-over" above
        ctx.composite = _COMPOSITE
        if _LN_FOLD and not needs_bwd and bn * t >= LN_FOLD_MIN_ROWS:
            return _block_forward_folded(x, x2d, bn, t, heads, ctx.grad_params, (w_in16, w_out16, w1_16, w2_16),
                                         float(p) if training else 0.0, float(attn_p) if training else 0.0, eps_msa, eps_mlp)
        if _COMPOSITE:
            x2, saved, drop, adrop = _block_forward_c(
                x2d, bn, t, heads, ctx.grad_params, (w_in16, w_out16, w1_16, w2_16), float(p) if training else 0.0,
                float(attn_p) if training else 0.0, eps_msa, eps_mlp, needs_bwd)
            if needs_bwd:
                ctx.save_for_backward(a_ln_w, m_ln_w, w_in16, w_out16, w1_16, w2_16, *saved)
            ctx.meta = (bn, t, heads, 0)
            ctx.drop = drop
            ctx.adrop = adrop
            out = x2.view(bn, t, d)
            if needs_bwd:
                _note_block_output(out, drop[0], drop[2])
            return out
        x1, s_msa, adrop = _msa_forward(x2d, bn, t, heads, a_ln_w, a_ln_b, w_in16, b_in, w_out16, b_out, x2d,
                                        float(attn_p) if training else 0.0, eps_msa)
        x2, s_mlp, drop = _mlp_forward(x1, m_ln_w, m_ln_b, w1_16, b1, w2_16, b2, float(p) if training else 0.0, x1,
                                       keep_preact=needs_bwd, eps=eps_mlp)
        if needs_bwd:
            ctx.save_for_backward(a_ln_w, m_ln_w, w_in16, w_out16, w1_16, w2_16, *s_msa, *s_mlp)
        ctx.meta = (bn, t, heads, len(s_msa))
        ctx.drop = drop
        ctx.adrop = adrop
        out = x2.view(bn, t, d)
        if needs_bwd:
            _note_block_output(out, drop[0], drop[2])
        return out

    @staticmethod
    def backward(ctx, g):
        a_ln_w, m_ln_w, w_in16, w_out16, w1_16, w2_16, *rest = ctx.saved_tensors
        bn, t, heads, n_msa = ctx.meta
        d = g.shape[-1]
        g2 = _f32c(g).view(bn * t, d)
        if ctx.composite:
            g0, gr = _block_backward_c(g2, tuple(rest), bn, t, heads, (a_ln_w, m_ln_w, w_in16, w_out16, w1_16, w2_16),
                                       ctx.drop, ctx.adrop, ctx.consumer_drop, ctx.grad_params)
            return (g0.view(bn, t, d), *gr, None, None, None, None, None, None, None, None, None, None, None)
        s_msa, s_mlp = tuple(rest[:n_msa]), tuple(rest[n_msa:])
        m = w1_16.shape[0]
        gp = ctx.grad_params
        direct = any(id(q) in _GRAD_DEST for q in gp)
        # one zero-fill for the five accumulate-into buffers -- unless they live in a registered destination
        # (False = no arena: whatever is not written to a destination is allocated on demand)
        arena = False if direct else _ZeroArena(g2, [(d, m), (m,), (m, d), (d, d), (3 * d, d)])   # in the order taken
        g1, (m_dgam, m_dbet, dw1, db1, dw2, db2), g1_16, db_out = _mlp_backward(
            g2, s_mlp, ctx.drop, m_ln_w, w1_16, w2_16, has_residual=True, want_bf16_out=True, arena=arena,
            params=gp[6:12], next_bias_param=gp[5])
        g0, (a_dgam, a_dbet, dw_in, db_in, dw_out, db_out) = _msa_backward(
            g1, s_msa, bn, t, heads, a_ln_w, w_in16, w_out16, has_residual=True, g16=g1_16, db_out=db_out,
            adrop=ctx.adrop, arena=arena, consumer_drop=ctx.consumer_drop, params=gp[0:6])
        return (g0.view(bn, t, d), a_dgam, a_dbet, dw_in, db_in, dw_out, db_out, m_dgam, m_dbet, dw1, db1, dw2, db2,
                None, None, None, None, None, None, None, None, None, None, None)


# ---------------------------------------------------------------------------------------------------------
# final LayerNorm (+ classifier head)  (models/vit.py:232-235 ; headless: models/vit_no_classifier.py:225-226)
# ---------------------------------------------------------------------------------------------------------
class HeadFn(torch.autograd.Function):
    """logits = Linear(LayerNorm(x)[:, 0]). Only the class-token row of each image is normalised: the reference
    normalises all T rows and discards T-1 of them; the result is identical."""

    @staticmethod
    def forward(ctx, x, ln_w, ln_b, w, b, eps: float = LN_EPS):
        _require_cuda(x, "ViT head")
        x = _f32c(x)
        bn, t, d = x.shape
        c = w.shape[0]
        xn = _empty((bn, d), torch.float32, x)
        mean = _empty((bn,), torch.float32, x)
        rstd = _empty((bn,), torch.float32, x)
        K.layernorm_fwd(x, t * d, ln_w, ln_b, None, xn, mean, rstd, bn, d, eps)
        logits = _empty((bn, c), torch.float32, x)
        K.head_fwd(xn, w, b, logits, bn, d, c)
        ctx.save_for_backward(x, ln_w, w, xn, mean, rstd)
        return logits

    @staticmethod
    def backward(ctx, gl):
        x, ln_w, w, xn, mean, rstd = ctx.saved_tensors
        bn, t, d = x.shape
        c = w.shape[0]
        gl = _f32c(gl)
        dxn = _empty((bn, d), torch.float32, x)
        dw = _empty((c, d), torch.float32, x)
        db = _empty((c,), torch.float32, x)
        K.head_bwd(xn, w, gl, dxn, dw, db, bn, d, c)
        gx = _zeros((bn, t, d), x)   # only class-token rows receive gradient
        dgam = _empty((d,), torch.float32, x)
        dbet = _empty((d,), torch.float32, x)
        K.layernorm_bwd(None, dxn, x, t * d, mean, rstd, ln_w, None, gx, t * d, None, dgam, dbet, None, bn, d)
        return gx, dgam, dbet, dw, db, None


class LayerNormFn(torch.autograd.Function):
    """Full-sequence nn.LayerNorm with f32 output (headless ViT, models/vit_no_classifier.py:225-226)."""

    @staticmethod
    def forward(ctx, x, w, b, eps: float = LN_EPS):
        _require_cuda(x, "LayerNorm")
        x = _f32c(x)
        d = x.shape[-1]
        rows = x.numel() // d
        y = torch.empty_like(x)
        mean = _empty((rows,), torch.float32, x)
        rstd = _empty((rows,), torch.float32, x)
        K.layernorm_fwd(x, d, w, b, None, y, mean, rstd, rows, d, eps)
        ctx.save_for_backward(x, w, mean, rstd)
        return y

    @staticmethod
    def backward(ctx, g):
        x, w, mean, rstd = ctx.saved_tensors
        d = x.shape[-1]
        rows = x.numel() // d
        g = _f32c(g)
        gx = torch.empty_like(x)
        dgam = _empty((d,), torch.float32, x)
        dbet = _empty((d,), torch.float32, x)
        K.layernorm_bwd(None, g, x, d, mean, rstd, w, None, gx, d, None, dgam, dbet, None, rows, d)
        return gx, dgam, dbet, None


# ---------------------------------------------------------------------------------------------------------
# fused softmax cross-entropy  (nn.CrossEntropyLoss(), engine.py:53)
# ---------------------------------------------------------------------------------------------------------
class CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, stats: Optional[dict]):
        _require_cuda(logits, "CrossEntropyLoss")
        logits = _f32c(logits)
        if logits.dim() != 2:
            raise ValueError("b200vit CrossEntropyLoss expects logits [batch, classes]")
        if target.dtype != torch.int64:
            target = target.long()
        target = target.contiguous()
        bn, c = logits.shape
        loss = _empty((1,), torch.float32, logits)
        need_grad = ctx.needs_input_grad[0]
        dl = _empty((bn, c), torch.float32, logits) if need_grad else None
        correct = torch.empty((1,), dtype=torch.int32, device=logits.device)
        K.softmax_ce(logits, target, loss, dl, correct, bn, c)
        if stats is not None:
            stats["correct"] = correct
            stats["batch"] = bn
        if need_grad:
            ctx.save_for_backward(dl)
        return loss.view(())

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        out = torch.empty_like(dl)
        K.scale_by_scalar(dl, _f32c(g).reshape(1), out)
        return out, None, None
