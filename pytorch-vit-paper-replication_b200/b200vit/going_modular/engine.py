"""train_step / test_step / train with the reference's signatures and return values
(reference: going_modular/going_modular/engine.py:9-79, 81-130, 132-211).

What differs from the reference is only *how* the same numbers are produced:
  * each batch's loss / accuracy goes to pinned host memory with a non-blocking copy and the host waits ONCE per
    epoch instead of two blocking `.item()` calls per batch (reference engine.py:54, 74, 121, 125) -- the returned
    floats are the same means of per-batch means;
  * the host->device copy of batch i+1 is issued on a side stream while batch i computes (pinned loaders overlap);
  * with `b200vit.optim.FusedAdam`, gradient clipping at global norm 1 (engine.py:63) is folded into the
    optimizer kernel; any other torch optimizer takes the reference's clip_grad_norm_ + step() path;
  * with `b200vit.nn.CrossEntropyLoss`, the accuracy count comes out of the loss kernel.
Any iterable of (X, y) batches with a len() works as a dataloader, as in the reference.

Two additions that the reference does not have, both inert unless asked for:
  * under `b200vit.dp.DataParallel` the two epoch metrics are averaged over the ranks (one 2-element all-reduce per
    epoch), so every rank returns the figure for the global batches;
  * `train_step(..., accumulation_steps=k)` (keyword-only) makes one optimizer step per k loader batches -- gradients
    of the k micro-batches are averaged, all but the last backward run under `DataParallel.no_sync()` -- which is how a
    global batch that does not fit (BASELINE.json configs[2]: 4096 images on 1 or 2 GPUs) is run with unchanged math.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import torch
from tqdm.auto import tqdm

from ..nn import CrossEntropyLoss as _FusedCE
from ..optim import FusedAdam as _FusedAdam

MAX_GRAD_NORM = 1.0  # Table 3 of the ViT paper; reference engine.py:63


def _device_batches(dataloader, device):
    """Yield (X, y) on `device`, with the host->device copy of the NEXT batch already in flight on a side stream
    while the caller computes on the current one (the reference copies synchronously inside the loop,
    engine.py:47). Batches that already live on the device pass through untouched."""
    dev = torch.device(device)
    if dev.type != "cuda":
        for X, y in dataloader:
            yield X.to(dev), y.to(dev)
        return
    copy_stream = torch.cuda.Stream(device=dev)
    compute = torch.cuda.current_stream(dev)

    def stage(batch):
        X, y = batch
        if X.device == dev and y.device == dev:
            return X, y, None
        with torch.cuda.stream(copy_stream):
            Xd, yd = X.to(dev, non_blocking=True), y.to(dev, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return Xd, yd, ev

    it = iter(dataloader)
    try:
        nxt = stage(next(it))
    except StopIteration:
        return
    while nxt is not None:
        Xd, yd, ev = nxt
        try:
            nxt = stage(next(it))
        except StopIteration:
            nxt = None
        if ev is not None:
            compute.wait_event(ev)
            Xd.record_stream(compute)
            yd.record_stream(compute)
        yield Xd, yd


class _EpochMetrics:
    """Per-batch (loss, accuracy) pairs. On CUDA every batch's pair is copied to pinned host memory right away
    with a non-blocking D2H copy (8 bytes per step) and the host waits only once, at the end of the epoch -- the
    reference blocks on two `.item()` calls per batch instead (engine.py:54, 74)."""

    def __init__(self, n: int, device):
        self.n = n
        self.dev = torch.device(device)
        self.i = 0
        if self.dev.type == "cuda":
            self.host = torch.zeros((max(n, 1), 2), dtype=torch.float32, pin_memory=True)
        else:
            self.host = torch.zeros((max(n, 1), 2), dtype=torch.float32)

    def add(self, loss: torch.Tensor, acc: torch.Tensor) -> None:
        pair = torch.stack((loss.detach().to(torch.float32).reshape(()), acc.reshape(())))
        if self.i >= self.host.shape[0]:          # loader longer than its len(): grow
            grown = torch.zeros((2 * self.host.shape[0], 2), dtype=torch.float32, pin_memory=self.host.is_pinned())
            if self.dev.type == "cuda":
                torch.cuda.current_stream(self.dev).synchronize()
            grown[:self.i] = self.host[:self.i]
            self.host = grown
        self.host[self.i].copy_(pair, non_blocking=True)
        self.i += 1

    def means(self, model=None) -> Tuple[float, float]:
        if self.dev.type == "cuda":
            torch.cuda.current_stream(self.dev).synchronize()
        tot = self.host[:self.i].double().sum(dim=0)
        loss, acc = float(tot[0]) / self.n, float(tot[1]) / self.n
        reduce = getattr(model, "reduce_metrics", None)        # b200vit.dp.DataParallel: mean over the ranks
        return reduce(loss, acc) if callable(reduce) else (loss, acc)


def _batch_correct(loss_fn, logits: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """Number of correct top-1 predictions as a device scalar (float)."""
    if isinstance(loss_fn, _FusedCE) and loss_fn.last_correct is not None:
        return loss_fn.last_correct.to(torch.float32).reshape(())
    return (logits.argmax(dim=1) == y).sum().to(torch.float32)


def train_step(model: torch.nn.Module,
               dataloader: torch.utils.data.DataLoader,
               loss_fn: torch.nn.Module,
               optimizer: torch.optim.Optimizer,
               lr_scheduler: torch.optim.lr_scheduler,
               device: torch.device,
               *, accumulation_steps: int = 1) -> Tuple[float, float]:
    """One epoch of training; returns (mean batch loss, mean batch accuracy)."""
    model.train()
    metrics = _EpochMetrics(len(dataloader), device)
    fused_opt = isinstance(optimizer, _FusedAdam)
    k = int(accumulation_steps)
    if k < 1:
        raise ValueError("accumulation_steps must be >= 1")
    no_sync = getattr(model, "no_sync", None)
    inv_k = None        # upstream gradient 1/k handed to backward(): averages the k micro-batch gradients

    for batch, (X, y) in enumerate(_device_batches(dataloader, device)):
        y_pred = model(X)
        loss = loss_fn(y_pred, y)
        metrics.add(loss, _batch_correct(loss_fn, y_pred.detach(), y) / len(y_pred))

        if k == 1:
            optimizer.zero_grad()
            loss.backward()
        else:
            micro = batch % k
            if micro == 0:
                optimizer.zero_grad()
            if inv_k is None:
                inv_k = torch.full_like(loss.detach(), 1.0 / k)
            if micro != k - 1:
                if callable(no_sync):
                    with no_sync():
                        loss.backward(inv_k)
                else:
                    loss.backward(inv_k)
                continue
            loss.backward(inv_k)

        if fused_opt:
            optimizer.step(max_norm=MAX_GRAD_NORM)
        else:
            torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=MAX_GRAD_NORM)
            optimizer.step()
        lr_scheduler.step()

    return metrics.means(model)


def test_step(model: torch.nn.Module,
              dataloader: torch.utils.data.DataLoader,
              loss_fn: torch.nn.Module,
              device: torch.device) -> Tuple[float, float]:
    """One evaluation pass; returns (mean batch loss, mean batch accuracy)."""
    model.eval()
    metrics = _EpochMetrics(len(dataloader), device)
    with torch.inference_mode():
        for batch, (X, y) in enumerate(_device_batches(dataloader, device)):
            logits = model(X)
            loss = loss_fn(logits, y)
            metrics.add(loss, _batch_correct(loss_fn, logits, y) / len(logits))
    return metrics.means(model)


test_step.__test__ = False  # the name comes from the reference API; it is not a pytest test


def train(model: torch.nn.Module,
          train_dataloader: torch.utils.data.DataLoader,
          test_dataloader: torch.utils.data.DataLoader,
          optimizer: torch.optim.Optimizer,
          loss_fn: torch.nn.Module,
          lr_scheduler: torch.optim.lr_scheduler,
          epochs: int,
          device: torch.device) -> Dict[str, List]:
    """Runs `epochs` x (train_step, test_step); prints and returns the four metric lists per epoch."""
    results = {"train_loss": [], "train_acc": [], "test_loss": [], "test_acc": []}
    model.to(device)
    for epoch in tqdm(range(epochs)):
        train_loss, train_acc = train_step(model=model, dataloader=train_dataloader, loss_fn=loss_fn,
                                           optimizer=optimizer, lr_scheduler=lr_scheduler, device=device)
        test_loss, test_acc = test_step(model=model, dataloader=test_dataloader, loss_fn=loss_fn, device=device)
        print(f"Epoch: {epoch+1} | train_loss: {train_loss:.4f} | train_acc: {train_acc:.4f} | "
              f"test_loss: {test_loss:.4f} | test_acc: {test_acc:.4f}")
        results["train_loss"].append(train_loss)
        results["train_acc"].append(train_acc)
        results["test_loss"].append(test_loss)
        results["test_acc"].append(test_acc)
    return results
