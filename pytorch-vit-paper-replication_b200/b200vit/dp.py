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
  * the flat buffers are all-reduced (SUM) asynchronously on a side stream -- each one as soon as it is complete
    (default), or all of them from the end-of-backward engine callback (`overlap=False`). On one NVLink / NVSwitch node
    (<= 8 ranks) the reduction is the library's own peer-memory kernel (csrc/peer_reduce.cu, b200vit_allreduce_peer_f32):
    the buckets live in a symmetric allocation every rank maps, and one 64-thread-per-CTA kernel without shared memory
    sums slice `rank` of all ranks' buckets and writes it back to all of them, sharing SMs with the backward kernels it
    overlaps (NCCL's all-reduce CTAs cannot, and stalled the persistent GEMMs: 1.4 ms of a 34.4 ms step on 2 GPUs).
    B200VIT_DP_ALLREDUCE=nccl, another backend, more than 8 ranks or a failed rendezvous fall back to ncclAllReduce;
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


def _padded(n: int) -> int:
    return (n + 63) // 64 * 64          # bucket starts stay 256-byte aligned inside one arena


class _PeerWork:
    """Handle of one peer-memory all-reduce in flight (same wait() contract as torch.distributed's Work)."""

    def __init__(self, done: "torch.cuda.Event"):
        self.done = done

    def wait(self) -> None:
        torch.cuda.current_stream().wait_event(self.done)       # stream-ordered, the host does not block


class _PeerTransport:
    """Symmetric arena for all gradient buckets + the flag words of b200vit_allreduce_peer_f32 (include/b200vit.h).
    torch.distributed._symmetric_memory does the mapping of the peers' memory (plumbing); the reduction itself is the
    library's kernel, launched on a side stream after an event recorded on the backward stream."""

    def __init__(self, floats: int, device: torch.device, group, world: int, rank: int):
        import torch.distributed._symmetric_memory as symm
        from . import _C
        gname = (group if group is not None else dist.group.WORLD).group_name
        self.arena = symm.empty(floats, dtype=torch.float32, device=device)
        self.flags = symm.empty(_C.PEER_FLAG_WORDS, dtype=torch.int32, device=device)
        self.arena.zero_()
        self.flags.zero_()
        self._h = symm.rendezvous(self.arena, gname)
        self._hf = symm.rendezvous(self.flags, gname)
        torch.cuda.synchronize(device)
        self._hf.barrier(channel=0)             # every rank's flag words are zero before anybody signals
        torch.cuda.synchronize(device)
        self.bufs = [int(p) for p in self._h.buffer_ptrs]
        self.flag_ptrs = [int(p) for p in self._hf.buffer_ptrs]
        if len(self.bufs) != world or self.bufs[rank] != self.arena.data_ptr():
            raise RuntimeError("symmetric memory handle does not describe this process group")
        self.world, self.rank, self.device = world, rank, device
        self.stream = torch.cuda.Stream(device=device)
        # twice the SM count: every SM gets a CTA wherever the hardware places them; the surplus ones leave at once
        self.ctas = int(os.environ.get("B200VIT_DP_PEER_CTAS", 2 * torch.cuda.get_device_properties(device).multi_processor_count))
        self.seq = 0
        import ctypes
        self._flag_table = (ctypes.c_void_p * world)(*self.flag_ptrs)
        self._ctypes = ctypes

    def make_events(self):
        """(ready, done) events of one bucket, reused step after step."""
        return torch.cuda.Event(), torch.cuda.Event()

    def all_reduce(self, flat: torch.Tensor, ready: "torch.cuda.Event", done: "torch.cuda.Event") -> _PeerWork:
        from . import _C
        off = flat.data_ptr() - self.arena.data_ptr()
        assert 0 <= off and off + flat.numel() * 4 <= self.arena.numel() * 4 and flat.numel() % 4 == 0
        ready.record()                                  # on the stream backward is running on
        self.stream.wait_event(ready)
        self.seq += 1
        table = (self._ctypes.c_void_p * self.world)(*[b + off for b in self.bufs])
        with torch.cuda.device(self.device):
            _C.check(_C.lib().b200vit_allreduce_peer_f32(table, self._flag_table, self.world, self.rank, flat.numel(),
                                                         self.seq & 0xFFFFFFFF, self.ctas, self.stream.cuda_stream),
                     "b200vit_allreduce_peer_f32")
        _C.launch_count += 1
        done.record(self.stream)
        return _PeerWork(done)


class _Bucket:
    def __init__(self, params: List[nn.Parameter], flat: Optional[torch.Tensor] = None):
        self.params = params
        n = sum(p.numel() for p in params)
        # `flat` (a slice of the symmetric arena, zero, length _padded(n)) when the peer transport is in use
        self.flat = flat if flat is not None else torch.zeros(n, dtype=params[0].dtype, device=params[0].device)
        self.ready = self.done = None
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
        groups: List[List[nn.Parameter]] = []
        cur, cur_bytes = [], 0
        for p in reversed(params):
            cur.append(p)
            cur_bytes += p.numel() * p.element_size()
            if cur_bytes >= limit:
                groups.append(cur)
                cur, cur_bytes = [], 0
        if cur:
            groups.append(cur)
        self.peer = self._make_peer_transport(groups)
        self.buckets: List[_Bucket] = []
        off = 0
        for g in groups:
            if self.peer is not None:
                n = _padded(sum(p.numel() for p in g))
                b = _Bucket(g, self.peer.arena[off:off + n])
                b.ready, b.done = self.peer.make_events()
                off += n
            else:
                b = _Bucket(g)
            self.buckets.append(b)
        self._where = {}
        for b in self.buckets:
            for i, p in enumerate(b.params):
                self._where[p] = (b, i)
                p.register_post_accumulate_grad_hook(self._on_grad)
                F_.register_grad_destination(p, lambda b=b, i=i: b.view(i))
        self._callback_queued = False

    def _make_peer_transport(self, groups) -> Optional[_PeerTransport]:
        """The peer-memory transport if every rank can have it (all ranks decide alike), else None (NCCL)."""
        want = os.environ.get("B200VIT_DP_ALLREDUCE", "peer")
        if want not in ("peer", "nccl"):
            raise ValueError(f"B200VIT_DP_ALLREDUCE={want!r}: expected 'peer' or 'nccl'")
        if self.world < 2 or dist.get_backend(self.group) != "nccl" or not groups:
            return None
        dev = groups[0][0].device
        peer, ok = None, 0
        if want == "peer" and self.world <= 8 and dev.type == "cuda" and all(p.dtype == torch.float32 and p.device == dev
                                                                             for g in groups for p in g):
            try:
                peer = _PeerTransport(sum(_padded(sum(p.numel() for p in g)) for g in groups), dev, self.group,
                                      self.world, self.rank)
                ok = 1
            except Exception as e:          # no symmetric memory in this torch / on this box
                if os.environ.get("B200VIT_DP_DEBUG"):
                    print(f"[b200vit.dp] rank {self.rank}: peer transport unavailable: {e!r}", flush=True)
        agree = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=self.group)
        return peer if int(agree.item()) == 1 else None

    def _all_reduce(self, b: _Bucket):
        if self.peer is not None:
            return self.peer.all_reduce(b.flat, b.ready, b.done)
        return dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.reduce_group, async_op=True)

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
            b.work = self._all_reduce(b)

    def _finish(self) -> None:
        for b in self.buckets:
            if b.work is None and b.pending != len(b.params):
                # not started yet (overlap off), or some parameters of this bucket got no gradient this step:
                # reduce what is there (the missing ones are zero)
                if b.pending != 0:
                    for j, p in enumerate(b.params):
                        if p.grad is None:
                            b.view(j).zero_()
                b.work = self._all_reduce(b)
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
