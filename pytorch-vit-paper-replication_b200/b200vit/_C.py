"""ctypes binding of libb200vit.so (see include/b200vit.h).

The library is the ONLY compute backend: if it is missing or a call fails, a RuntimeError is raised.
There is no PyTorch / CPU fallback behind these functions.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libb200vit.so"
ABI_VERSION = 1

_p = C.c_void_p
_i = C.c_int
_ll = C.c_longlong
_ull = C.c_ulonglong
_f = C.c_float
_u = C.c_uint

# name -> argtypes (return type is always int unless noted)
_SIGNATURES = {
    "b200vit_gemm": [_i, _i, _i, _i, _i, _p, _ll, _p, _ll, _p, _ll, _p, _ll, _p, _p, _ll, _f, _ull, _ull, _i, _i, _p],
    "b200vit_cast_f32_to_bf16": [_p, _p, _ll, _p],
    "b200vit_im2col_patches": [_p, _p, _i, _i, _i, _i, _i, _p],
    "b200vit_im2col_patches_u8": [_p, _p, _i, _i, _i, _i, _i, _p],
    "b200vit_class_token_rows": [_p, _p, _p, _i, _i, _i, _f, _ull, _ull, _p],
    "b200vit_embed_param_grads": [_p, _p, _p, _p, _i, _i, _i, _f, _ull, _ull, _p],
    "b200vit_patch_grad_rows": [_p, _p, _i, _i, _i, _f, _ull, _ull, _p],
    "b200vit_layernorm_fwd": [_p, _ll, _p, _p, _p, _p, _p, _p, _ll, _i, _f, _p],
    "b200vit_layernorm_bwd": [_p, _p, _p, _ll, _p, _p, _p, _p, _p, _ll, _p, _p, _p, _p, _ll, _i, _f, _ull, _ull, _p],
    "b200vit_colsum_bf16": [_p, _ll, _p, _ll, _i, _p],
    "b200vit_dropout_cast": [_p, _p, _p, _ll, _i, _f, _ull, _ull, _p],
    "b200vit_attention_fwd": [_p, _p, _p, _i, _i, _i, _i, _f, _p],
    "b200vit_attention_bwd": [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _f, _p],
    "b200vit_attention_bwd_bias": [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _f, _p],
    "b200vit_attention_fwd_dropout": [_p, _p, _p, _i, _i, _i, _i, _f, _f, _ull, _ull, _p],
    "b200vit_attention_bwd_dropout": [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _f, _f, _ull, _ull, _p],
    "b200vit_head_fwd": [_p, _p, _p, _p, _i, _i, _i, _p],
    "b200vit_head_bwd": [_p, _p, _p, _p, _p, _p, _i, _i, _i, _p],
    "b200vit_softmax_ce": [_p, _p, _p, _p, _p, _i, _i, _p],
    "b200vit_scale_by_scalar": [_p, _p, _p, _ll, _p],
    "b200vit_softmax_top1": [_p, _p, _p, _i, _i, _p],
    "b200vit_l2norm_sq": [_p, _p, _i, _ll, _p, _i, _p, _p],
    "b200vit_l2norm_sq_noalloc": [_p, _p, _i, _ll, _p, _p],
    "b200vit_adam_step": [_p, _i, _ll, _p, _p, _i, _f, _f, _f, _i, _i, _p, _f, _f, _p],
    "b200vit_set_sm_reserve": [_i],
    "b200vit_set_gemm_schedule": [_i],
    "b200vit_allreduce_peer_f32": [_p, _p, _i, _i, _ll, _u, _i, _p],
    "b200vit_gemm_ln": [_i, _i, _i, _i, _i, _p, _ll, _p, _ll, _p, _ll, _p, _ll, _p, _p, _ll, _f, _ull, _ull, _i, _p, _ll, _p,
                        _p, _p, _i, _f, _p],
    "b200vit_fold_ln_weight": [_p, _p, _p, _p, _p, _p, _p, _i, _i, _p],
    "b200vit_class_token_rows_ln": [_p, _p, _p, _p, _p, _i, _i, _i, _f, _ull, _ull, _p],
    "b200vit_encoder_block_fwd": [_p, _p],
    "b200vit_encoder_block_bwd": [_p, _p],
    "b200vit_split_bf16x3": [_p, _ll, _ll, _i, _p, _ll, _ll, _i, _f, _ull, _ull, _i, _p],
    "b200vit_im2col_patches_f32": [_p, _p, _i, _i, _i, _i, _i, _p],
    "b200vit_gelu_f32_fwd": [_p, _p, _ll, _f, _ull, _ull, _p],
    "b200vit_gelu_f32_bwd": [_p, _p, _p, _ll, _f, _ull, _ull, _p],
    "b200vit_colsum_f32": [_p, _ll, _ll, _i, _p, _f, _ull, _ull, _i, _p],
    "b200vit_attention_f32_fwd": [_p, _p, _p, _i, _i, _i, _i, _f, _p],
    "b200vit_attention_f32_bwd": [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _f, _p],
}

# GEMM enums (include/b200vit.h)
OP_NT, OP_NN, OP_TN = 0, 1, 2
EPI_BF16, EPI_BIAS_GELU, EPI_BIAS_RES, EPI_DGELU, EPI_ACCUM, EPI_F32, EPI_PATCH = range(7)
EPI_RES_STATS, EPI_PATCH_STATS, EPI_LN_BF16, EPI_LN_GELU = 7, 8, 9, 10      # LayerNorm fold (b200vit_gemm_ln)


class BlockFwdArgs(C.Structure):
    """b200vit_block_fwd_args (include/b200vit.h)."""
    _fields_ = ([("struct_bytes", _i), ("batch", _i), ("tokens", _i), ("heads", _i), ("dim", _i), ("mlp", _i),
                 ("eps_msa", _f), ("eps_mlp", _f), ("mlp_p", _f), ("attn_p", _f),
                 ("key1_seed", _ull), ("key1_offset", _ull), ("key2_seed", _ull), ("key2_offset", _ull),
                 ("attn_seed", _ull), ("attn_offset", _ull)] +
                [(n, _p) for n in ("x", "ln1_w", "ln1_b", "b_in", "b_out", "ln2_w", "ln2_b", "b1", "b2",
                                   "w_in16", "w_out16", "w1_16", "w2_16",
                                   "ln1", "qkv", "o", "ln2", "a", "h",
                                   "mean1", "rstd1", "lse", "x1", "mean2", "rstd2", "x2")])


class BlockBwdArgs(C.Structure):
    """b200vit_block_bwd_args (include/b200vit.h)."""
    _fields_ = ([("struct_bytes", _i), ("batch", _i), ("tokens", _i), ("heads", _i), ("dim", _i), ("mlp", _i),
                 ("mlp_p", _f), ("attn_p", _f),
                 ("key1_seed", _ull), ("key1_offset", _ull), ("key2_seed", _ull), ("key2_offset", _ull),
                 ("attn_seed", _ull), ("attn_offset", _ull),
                 ("g", _p), ("g16_ready", _i), ("zero_accumulators", _i)] +
                [(n, _p) for n in ("x", "mean1", "rstd1", "lse", "x1", "mean2", "rstd2",
                                   "ln1", "qkv", "o", "ln2", "a", "h",
                                   "ln1_w", "ln2_w", "w_in16", "w_out16", "w1_16", "w2_16",
                                   "g16", "da", "dout", "dqkv", "g1_16", "g1", "delta",
                                   "gx",
                                   "d_ln1_w", "d_ln1_b", "dw_in", "db_in", "dw_out", "db_out", "d_ln2_w", "d_ln2_b",
                                   "dw1", "db1", "dw2", "db2",
                                   "gx16_out", "dbn_out")] +
                [("consumer_p", _f), ("consumer_seed", _ull), ("consumer_offset", _ull)])


class AdamTensor(C.Structure):
    """Record of b200vit_adam_step's tensor table (must match csrc/optim.cu)."""
    _fields_ = [("p", _p), ("g", _p), ("m", _p), ("v", _p), ("shadow", _p), ("n", _ll), ("group", _i), ("pad", _i)]


PEER_FLAG_WORDS = 320       # B200VIT_PEER_FLAG_WORDS of include/b200vit.h
_lib = None
_override = None  # a test double may be installed with set_library(); never done by product code


def library_path() -> Path:
    return Path(os.environ.get("B200VIT_LIB", str(_LIB_PATH)))


def set_library(obj) -> None:
    """Install an object exposing the C-ABI function names (used by tests/ to inject a CPU test double for
    host-logic tests; product code never calls this)."""
    global _override
    _override = obj


def lib():
    """Load libb200vit.so (once) and return the ctypes handle; raise loudly if it is not there."""
    global _lib
    if _override is not None:
        return _override
    if _lib is None:
        path = library_path()
        if not path.exists():
            raise RuntimeError(
                f"libb200vit.so not found at {path}. Build it with "
                f"`python pytorch-vit-paper-replication_b200/build.py` (or __graft_entry__.build()). "
                f"b200vit has no CPU / PyTorch fallback."
            )
        handle = C.CDLL(str(path))
        handle.b200vit_abi_version.restype = _i
        handle.b200vit_abi_version.argtypes = []
        handle.b200vit_last_error.restype = C.c_char_p
        handle.b200vit_last_error.argtypes = []
        got = handle.b200vit_abi_version()
        if got != ABI_VERSION:
            raise RuntimeError(f"libb200vit.so ABI version {got} != expected {ABI_VERSION}; rebuild the library")
        for name, argtypes in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = _i
            fn.argtypes = argtypes
        handle.b200vit_tmap_cache_stats.restype = None
        handle.b200vit_tmap_cache_stats.argtypes = [C.POINTER(_ll), C.POINTER(_ll)]
        _lib = handle
    return _lib


def exported_symbols():
    return ["b200vit_abi_version", "b200vit_last_error", "b200vit_tmap_cache_stats", *_SIGNATURES.keys()]


def set_sm_reserve(sms: int) -> None:
    """Keep `sms` SMs out of every grid launched from now on (see include/b200vit.h)."""
    check(lib().b200vit_set_sm_reserve(int(sms)), "b200vit_set_sm_reserve")


def set_gemm_schedule(dynamic: bool) -> None:
    """Dynamic (claimed from a device counter) or static distribution of the 2-CTA GEMM's work items (include/b200vit.h)."""
    check(lib().b200vit_set_gemm_schedule(1 if dynamic else 0), "b200vit_set_gemm_schedule")


def tmap_cache_stats():
    """(hits, misses) of the library's tensor-map cache."""
    h, m = _ll(0), _ll(0)
    lib().b200vit_tmap_cache_stats(C.byref(h), C.byref(m))
    return h.value, m.value


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().b200vit_last_error()
        if isinstance(msg, bytes):
            msg = msg.decode(errors="replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


# CUDA kernels launched by one successful call of each entry point (memsets are not counted).
KERNELS_PER_CALL = {"b200vit_l2norm_sq": 2, "b200vit_l2norm_sq_noalloc": 2,
                    "b200vit_head_bwd": 2, "b200vit_attention_f32_bwd": 2,
                    "b200vit_encoder_block_fwd": 7, "b200vit_encoder_block_bwd": 12}
launch_count = 0
call_count = 0          # Python -> C transitions (one per entry-point call)
want_device = None      # CUDA device ordinal of the tensors of the call being prepared (set by kernels.stream_ptr)
_current_device = None  # bound lazily to torch.cuda.current_device


_timing = None          # list of (name, first five args, start event, end event) while a CallTimer is active


class CallTimer:
    """Measurement aid: CUDA events (on the launching stream) around every entry-point call made inside the `with`
    block, so that the time each kernel family takes INSIDE a real step can be read back -- `by_name()` after the block
    synchronises and returns {entry point: [milliseconds per call, ...]} and `gemms()` the (op, epilogue, M, N, K, ms)
    list. The events sit between consecutive launches, which switches off the programmatic-dependent-launch overlap of
    a kernel's prologue with its predecessor's tail: the figures are a few microseconds per launch pessimistic."""

    def __enter__(self):
        global _timing
        self.records = _timing = []
        return self

    def __exit__(self, *a):
        global _timing
        _timing = None

    def by_name(self):
        import torch
        torch.cuda.synchronize()
        out = {}
        for name, _head, e0, e1 in self.records:
            out.setdefault(name, []).append(e0.elapsed_time(e1))
        return out

    def gemms(self):
        import torch
        torch.cuda.synchronize()
        return [(*head, e0.elapsed_time(e1)) for name, head, e0, e1 in self.records if name == "b200vit_gemm"]


def _timed_call(fn, name, args):
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    rc = fn(*args)
    e1.record()
    _timing.append((name, tuple(args[:5]), e0, e1))
    return rc


def call(name: str, *args) -> None:
    """Invoke one entry point and raise on a non-zero return. Kernels launch on the stream that kernels.stream_ptr()
    looked up for the tensors' device; if that is not the CUDA current device (a model on cuda:1 without a prior
    torch.cuda.set_device(1)) the call runs under a device guard, as torch's own operators do."""
    global launch_count, call_count, _current_device
    fn = getattr(lib(), name)
    dev = want_device
    if dev is not None:
        if _current_device is None:
            import torch
            _current_device = torch.cuda.current_device
        if dev != _current_device():
            import torch
            with torch.cuda.device(dev):
                check(fn(*args), name)
            launch_count += KERNELS_PER_CALL.get(name, 1)
            call_count += 1
            return
    check(fn(*args) if _timing is None else _timed_call(fn, name, args), name)
    launch_count += KERNELS_PER_CALL.get(name, 1)
    call_count += 1
