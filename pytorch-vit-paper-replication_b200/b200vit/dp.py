"""Plain data parallelism for the training step: one process per GPU, full replica per rank, one bucketed
gradient all-reduce per optimizer step, overlapped with the rest of backward (or, optionally, issued at its end).

The reference has no distributed code at all (single `model.to(device)`, going_modular/engine.py:180); this
wrapper is the one exchange step that data-parallel training of its hot path needs. It is module-agnostic:

    model = DataParallel(ViT(...).to(dev))          # broadcasts rank 0's weights
    opt = FusedAdam(model.parameters(), ...); opt.grad_scale = model.grad_scale   # = 1/world (sum -> mean)
    engine.train_step(model, loader, loss_fn, opt, sched, dev)                    # unchanged

Mechanics
  * parameters are grouped, in reverse registration order (~ the order backward produces gradients), into
    buckets of ~`bucket_mb` MB backed by one flat fp32 buffer each;
  * the b200vit encoder blocks write their parameter gradients STRAIGHT into their bucket slices (the weight-gradient
    GEMMs accumulate there; b200vit.functional.grad_out), so `p.grad` is born as a view of the bucket and nothing is
    copied; gradients produced by anything else (other modules, stock torch layers) are moved into their slice by a
    post-accumulate-grad hook, after which `p.grad` is that view too (the reduced values are seen by
    clip_grad_norm_ / any optimizer without a copy back);
  * the flat buffers are all-reduced (SUM) asynchronously on NCCL's own stream -- each one as soon as it is complete
    (default), or all of them from the end-of-backward engine callback (`overlap=False`);
  * the end-of-backward callback waits for the outstanding reductions, so gradients are final when
    `loss.backward()` returns (the reference engine calls clip_grad_norm_ right after, engine.py:60-63).
Gradients hold the SUM over ranks; `grad_scale` (1/world) is applied by FusedAdam on the fly, or call
`average_gradients()` before a stock optimizer.

Gradient accumulation over micro-batches: run every micro-step but the last under `no_sync()`. A second synchronised
backward on top of already reduced gradients (no `zero_grad()` and no `no_sync()` in between) would reduce the first
step's sum a second time; it is detected and refused.

Every rank gets its own family of dropout masks (functional.set_dropout_stream(rank)): ranks seeded alike -- the usual
`torch.manual_seed(42)` on all of them -- still drop independently over their shards of the global batch.
"""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist
from torch import nn

from . import functional as F_


class _Bucket:
    def __init__(self, params: List[nn.Parameter]):
        self.params = params
        n = sum(p.numel() for p in params)
        self.flat = torch.zeros(n, dtype=params[0].dtype, device=params[0].device)
        self.offsets = []
        off = 0
        for p in params:
            self.offsets.append(off)
            off += p.numel()
        self.pending = len(params)
        self.work = None

    def view(self, i: int) -> torch.Tensor:
        """A FRESH tensor object over parameter i's slice (autograd adopts a gradient without copying only when nobody
        else holds a reference to the tensor object it is handed)."""
        p = self.params[i]
        return self.flat[self.offsets[i]:self.offsets[i] + p.numel()].view(p.shape)


class DataParallel(nn.Module):
    def __init__(self, module: nn.Module, process_group=None, bucket_mb: float = 25.0, broadcast: bool = True,
                 overlap: Optional[bool] = None):
        super().__init__()
        # overlap=True (default) starts each bucket's all-reduce as soon as the bucket is complete, while backward
        # continues; overlap=False starts them all at the end of backward. Measured on 2 B200 (batch 256 per GPU):
        # 37.05 ms/step overlapped, 37.39 ms at the end, 36.1 ms single-GPU. B200VIT_DP_OVERLAP=0/1 overrides.
        if overlap is None:
            overlap = os.environ.get("B200VIT_DP_OVERLAP", "1") == "1"
        if "B200VIT_DP_BUCKET_MB" in os.environ:
            bucket_mb = float(os.environ["B200VIT_DP_BUCKET_MB"])
        self.overlap = bool(overlap)
        if not dist.is_initialized():
            raise RuntimeError("DataParallel needs torch.distributed to be initialised (one process per GPU)")
        self.module = module
        self.group = process_group
        self.world = dist.get_world_size(process_group)
        self.rank = dist.get_rank(process_group)
        self.grad_scale = 1.0 / self.world
        self.sync = True
        # Optional: SMs kept free for the all-reduce kernels while backward runs, and the CTA budget NCCL gets for them
        # (see _reserve()). OFF by default: measured on 2 B200 (256 images per GPU) 36.7-37.5 ms per step without it,
        # 36.8-37.3 ms with 4 SMs reserved, 37.8-38.0 ms with 2 or 8 -- no gain beyond the box-to-box spread.
        # B200VIT_DP_SM_RESERVE=k switches it on.
        self.sm_reserve = int(os.environ.get("B200VIT_DP_SM_RESERVE", "0")) if self.overlap else 0
        self._reserved = False
        self.reduce_group = process_group
        if self.sm_reserve > 0 and self.world > 1 and dist.get_backend(process_group) == "nccl":
            try:
                opts = dist.ProcessGroupNCCL.Options()
                opts.config.max_ctas = self.sm_reserve
                opts.config.min_ctas = 1
                ranks = dist.get_process_group_ranks(process_group) if process_group is not None else None
                self.reduce_group = dist.new_group(ranks=ranks, backend="nccl", pg_options=opts)
            except Exception:       # no such option in this torch / NCCL: fall back to sharing SMs
                self.sm_reserve = 0
                self.reduce_group = process_group
        else:
            self.sm_reserve = 0
        F_.set_dropout_stream(self.rank)
        if broadcast:
            src = dist.get_global_rank(process_group, 0) if process_group else 0
            with torch.no_grad():       # in place on the tensor itself: bumps its version, so bf16 shadows made before
                for t in list(module.parameters()) + list(module.buffers()):      # wrapping are re-cast, not stale
                    dist.broadcast(t, src=src, group=process_group)
        params = [p for p in module.parameters() if p.requires_grad]
        limit = int(bucket_mb * 1024 * 1024)
        self.buckets: List[_Bucket] = []
        cur, cur_bytes = [], 0
        for p in reversed(params):
            cur.append(p)
            cur_bytes += p.numel() * p.element_size()
            if cur_bytes >= limit:
                self.buckets.append(_Bucket(cur))
                cur, cur_bytes = [], 0
        if cur:
            self.buckets.append(_Bucket(cur))
        self._where = {}
        for b in self.buckets:
            for i, p in enumerate(b.params):
                self._where[p] = (b, i)
                p.register_post_accumulate_grad_hook(self._on_grad)
                F_.register_grad_destination(p, lambda b=b, i=i: b.view(i))
        self._callback_queued = False

    # -- hooks ---------------------------------------------------------------------------------------
    def _on_grad(self, p: nn.Parameter) -> None:
        b, i = self._where[p]
        g = p.grad
        # same tensor object as when it was reduced, and exactly one in-place write since: autograd's `grad += new`
        # (zero_grad(set_to_none=False) writes every slice of the bucket first; set_to_none=True makes new objects)
        if self.sync and getattr(g, "_b200vit_reduced", None) == g._version - 1:
            raise RuntimeError(
                "DataParallel: a synchronised backward accumulated into gradients that were already all-reduced; the "
                "earlier sum would be reduced again. Call zero_grad() between optimizer steps, and run every "
                "micro-batch but the last of a gradient-accumulation step under DataParallel.no_sync().")
        if g.data_ptr() != b.flat.data_ptr() + b.offsets[i] * b.flat.element_size():
            view = b.view(i)
            view.copy_(g)
            p.grad = view
        if not self.sync:
            return
        if not self._callback_queued:
            torch.autograd.Variable._execution_engine.queue_callback(self._finish)
            self._callback_queued = True
            self._reserve(True)
        b.pending -= 1
        if b.pending == 0 and self.overlap:
            b.work = dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.reduce_group, async_op=True)

    def _finish(self) -> None:
        for b in self.buckets:
            if b.work is None and b.pending != len(b.params):
                # not started yet (overlap off), or some parameters of this bucket got no gradient this step:
                # reduce what is there (the missing ones are zero)
                if b.pending != 0:
                    for j, p in enumerate(b.params):
                        if p.grad is None:
                            b.view(j).zero_()
                b.work = dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.reduce_group, async_op=True)
        self._reserve(False)
        for b in self.buckets:
            if b.work is not None:
                b.work.wait()
                b.work = None
                for p in b.params:
                    if p.grad is not None:
                        p.grad._b200vit_reduced = p.grad._version
            b.pending = len(b.params)
        self._callback_queued = False

    def _reserve(self, on: bool) -> None:
        """The compute kernels are persistent grids of one CTA per SM with a static split of the work, so an all-reduce
        kernel that takes SMs away can stall whole CTAs of the kernels it overlaps. With `sm_reserve` > 0 the library
        launches its grids that many SMs short from the first gradient of a synchronised backward until its end, and
        NCCL's communicator for the gradient buckets is capped at that many CTAs. (Did not pay on 2 GPUs -- see
        __init__ -- and is therefore off unless asked for.)"""
        if self.sm_reserve > 0 and on != self._reserved:
            from . import _C
            _C.set_sm_reserve(self.sm_reserve if on else 0)
            self._reserved = on

    # -- API -----------------------------------------------------------------------------------------
    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def average_gradients(self) -> None:
        """Turn the summed gradients into means (only needed with optimizers that have no grad_scale)."""
        for b in self.buckets:
            b.flat.mul_(self.grad_scale)

    def reduce_metrics(self, loss: float, acc: float) -> Tuple[float, float]:
        """Mean over ranks of the two epoch metrics (one 2-element all-reduce per epoch): what engine.train_step /
        test_step return is then the figure for the GLOBAL batches, as a single-process run of the reference over the
        whole data would report (engine.py:77-79, 128-130), identical on every rank."""
        dev = self.buckets[0].flat.device if self.buckets else torch.device("cpu")
        t = torch.tensor([loss, acc], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        t /= self.world
        return float(t[0]), float(t[1])

    class _NoSync:
        def __init__(self, dp):
            self.dp = dp

        def __enter__(self):
            self.prev = self.dp.sync
            self.dp.sync = False

        def __exit__(self, *a):
            self.dp.sync = self.prev

    def no_sync(self):
        """Context manager: accumulate gradients locally without communication (gradient accumulation)."""
        return DataParallel._NoSync(self)
