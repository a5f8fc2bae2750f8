"""FusedAdam: torch.optim.Adam semantics in one multi-tensor CUDA launch.

Drop-in for the optimizer the reference notebook builds (`torch.optim.Adam([{decay, weight_decay: 0.03},
{no_decay, weight_decay: 0.0}], lr=1e-3, betas=(0.9, 0.999))`, 08_pytorch_paper_replicating_self.ipynb
:2818-2823) and that `engine.train_step` steps at going_modular/engine.py:66. It is a real
torch.optim.Optimizer: `param_groups` are honoured every step, so LinearLR / SequentialLR schedulers that
rewrite `group['lr']` (engine.py:68) work unchanged, and `state_dict()` uses Adam's keys
(step / exp_avg / exp_avg_sq).

Extras over torch.optim.Adam, all optional:
  * `step(max_norm=1.0)` folds torch.nn.utils.clip_grad_norm_ (engine.py:63) into the update: one
    multi-tensor sum-of-squares launch, then the clip coefficient is applied inside the Adam kernel.
  * `grad_scale` multiplies gradients on the fly (1/world_size for summed data-parallel gradients).
  * bf16 shadows of the GEMM weights (b200vit.functional.bf16_shadow) are refreshed in the same pass.
"""
from __future__ import annotations

import ctypes
import math
from typing import Iterable, List, Optional

import torch

from . import _C
from . import functional as F_
from . import kernels as K

_CHUNK = 8192  # elements per block in csrc/optim.cu (MT_CHUNK)


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0,
                 decoupled_weight_decay: bool = False):
        if lr < 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid betas: {betas}")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                        decoupled_weight_decay=decoupled_weight_decay)
        super().__init__(params, defaults)
        if len(self.param_groups) > 8:
            raise ValueError("FusedAdam supports at most 8 param groups")
        g0 = self.param_groups[0]
        for g in self.param_groups:
            if g["betas"] != g0["betas"] or g["eps"] != g0["eps"] or \
                    g["decoupled_weight_decay"] != g0["decoupled_weight_decay"]:
                raise ValueError("FusedAdam: betas / eps / decoupled_weight_decay must be the same in every group")
        self._table_key = None
        self._upload_ring = None
        self._upload_idx = 0
        self._table_dev = None
        self._ptrs_dev = None
        self._sizes_dev = None
        self._partials = None
        self._sumsq = None
        self._max_size = 0
        self._num = 0
        self.grad_scale = 1.0
        self._shared_step = None

    # -- tensor tables -------------------------------------------------------------------------------
    def _active(self):
        items = []
        for gi, group in enumerate(self.param_groups):
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                items.append((gi, p))
        return items

    def _ensure_state(self, p):
        st = self.state[p]
        if len(st) == 0:
            # parameters whose state is created before the first step share ONE step tensor (advanced once per step
            # instead of once per parameter); a parameter that joins later gets its own, as in torch.optim.Adam
            if self._shared_step is None:
                self._shared_step = torch.tensor(0.0, dtype=torch.float32)
            if float(self._shared_step) == 0.0:
                st["step"] = self._shared_step
            else:
                st["step"] = torch.tensor(0.0, dtype=torch.float32)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    def _build_tables(self, items):
        key = []
        for gi, p in items:
            st = self._ensure_state(p)
            sh = F_.shadow_entry(p)
            key.append((gi, p.data_ptr(), p.grad.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(),
                        sh.tensor.data_ptr() if sh is not None else 0, p.numel()))
        key = tuple(key)
        if key == self._table_key:
            return
        dev = items[0][1].device
        n = len(items)
        arr = (_C.AdamTensor * n)()
        ptrs = (ctypes.c_void_p * n)()
        sizes = (ctypes.c_longlong * n)()
        for i, (gi, pp, gp, mp, vp, sp, numel) in enumerate(key):
            arr[i].p, arr[i].g, arr[i].m, arr[i].v = pp, gp, mp, vp
            arr[i].shadow = sp if sp else None
            arr[i].n, arr[i].group, arr[i].pad = numel, gi, 0
            ptrs[i] = gp
            sizes[i] = numel
        for (gi, p) in items:
            if p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                raise RuntimeError("FusedAdam expects fp32 parameters and gradients")
            if not p.is_contiguous() or not p.grad.is_contiguous():
                raise RuntimeError("FusedAdam expects contiguous parameters and gradients")

        # Gradients are fresh tensors every step, so the table usually changes every step. The upload must not
        # synchronise the device (a blocking copy from pageable memory would drain the whole launch queue once per
        # step): stage through a small ring of pinned host buffers and copy asynchronously on the current stream.
        blob = bytes(arr) + bytes(ptrs) + bytes(sizes)
        nbytes = len(blob)
        ring = self._upload_ring
        if ring is None or ring[0][0].numel() < nbytes or ring[0][1].device != dev:
            cap = max(nbytes, 1 << 16)
            pin = dev.type == "cuda"
            ring = self._upload_ring = [
                (torch.empty(cap, dtype=torch.uint8, pin_memory=pin), torch.empty(cap, dtype=torch.uint8, device=dev),
                 [None]) for _ in range(4)]
            self._upload_idx = 0
        host, devbuf, ev = ring[self._upload_idx]
        self._upload_idx = (self._upload_idx + 1) % len(ring)
        if ev[0] is not None:
            ev[0].synchronize()        # four optimizer steps old: never blocks in practice
        ctypes.memmove(host.data_ptr(), blob, nbytes)
        devbuf[:nbytes].copy_(host[:nbytes], non_blocking=True)
        if dev.type == "cuda":
            ev[0] = torch.cuda.Event()
            ev[0].record()
        o1, o2 = ctypes.sizeof(arr), ctypes.sizeof(arr) + ctypes.sizeof(ptrs)
        self._table_dev = devbuf[:o1]
        self._ptrs_dev = devbuf[o1:o2]
        self._sizes_dev = devbuf[o2:nbytes]
        self._max_size = max(k[-1] for k in key)
        self._num = n
        chunks = (self._max_size + _CHUNK - 1) // _CHUNK
        need = chunks * n
        if self._partials is None or self._partials.numel() < need or self._partials.device != dev:
            self._partials = torch.empty(need, dtype=torch.float32, device=dev)
        if self._sumsq is None or self._sumsq.device != dev:
            self._sumsq = torch.empty(1, dtype=torch.float32, device=dev)
        self._table_key = key

    # -- public ---------------------------------------------------------------------------------------
    def grad_sumsq(self) -> torch.Tensor:
        """Device scalar: sum of squares of all current gradients (one launch; no host sync)."""
        items = self._active()
        if not items:
            raise RuntimeError("no gradients")
        self._build_tables(items)
        K.l2norm_sq(self._ptrs_dev, self._sizes_dev, self._num, self._max_size, self._partials, self._sumsq)
        return self._sumsq

    @torch.no_grad()
    def step(self, closure=None, max_norm: Optional[float] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        items = self._active()
        if not items:
            return loss
        self._build_tables(items)
        sumsq = None
        if max_norm is not None and max_norm > 0:
            K.l2norm_sq(self._ptrs_dev, self._sizes_dev, self._num, self._max_size, self._partials, self._sumsq)
            sumsq = self._sumsq
        g0 = self.param_groups[0]
        st0 = self.state[items[0][1]]
        step = int(st0["step"].item() if isinstance(st0["step"], torch.Tensor) else st0["step"]) + 1
        K.adam_step(self._table_dev, self._num, self._max_size, [g["lr"] for g in self.param_groups],
                    [g["weight_decay"] for g in self.param_groups], g0["betas"][0], g0["betas"][1], g0["eps"], step,
                    g0["decoupled_weight_decay"], sumsq, max_norm if max_norm else 0.0, self.grad_scale)
        # Advance every DISTINCT step tensor once. Parameters created together share one tensor object, and
        # state_dict() / load_state_dict() (deepcopy with a memo, also through torch.save / torch.load) preserve that
        # aliasing without restoring `_shared_step` -- so identity, not `_shared_step`, decides what gets bumped.
        seen = set()
        for _, p in items:
            st_step = self.state[p]["step"]
            if id(st_step) not in seen:
                seen.add(id(st_step))
                st_step += 1
        torch.autograd.graph.increment_version([p for _, p in items])   # the kernel wrote p behind autograd's back
        for _, p in items:
            sh = F_.shadow_entry(p)
            if sh is not None:
                sh.version = p._version                 # ... and refreshed its bf16 shadow in the same pass
        return loss

    def last_grad_norm(self) -> float:
        """Global gradient L2 norm used by the last clipped step (forces a device sync; for logging/tests)."""
        return math.sqrt(float(self._sumsq.item())) * self.grad_scale
