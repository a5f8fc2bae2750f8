"""train_step / test_step / train with the reference's signatures and return values
(reference: going_modular/going_modular/engine.py:9-79, 81-130, 132-211).

What differs from the reference is only *how* the same numbers are produced:
  * metrics are accumulated on the device and read back ONCE per epoch instead of two blocking `.item()`
    calls per batch (reference engine.py:54, 74, 121, 125) -- the returned floats are the same means of
    per-batch means;
  * host->device copies of the batch are issued non-blocking (pinned loaders overlap with compute);
  * with `b200vit.optim.FusedAdam`, gradient clipping at global norm 1 (engine.py:63) is folded into the
    optimizer kernel; any other torch optimizer takes the reference's clip_grad_norm_ + step() path;
  * with `b200vit.nn.CrossEntropyLoss`, the accuracy count comes out of the loss kernel.
Any iterable of (X, y) batches with a len() works as a dataloader, as in the reference.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import torch
from tqdm.auto import tqdm

from ..nn import CrossEntropyLoss as _FusedCE
from ..optim import FusedAdam as _FusedAdam

MAX_GRAD_NORM = 1.0  # Table 3 of the ViT paper; reference engine.py:63


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
               device: torch.device) -> Tuple[float, float]:
    """One epoch of training; returns (mean batch loss, mean batch accuracy)."""
    model.train()
    loss_sum = torch.zeros((), dtype=torch.float32, device=device)
    acc_sum = torch.zeros((), dtype=torch.float32, device=device)
    fused_opt = isinstance(optimizer, _FusedAdam)

    for batch, (X, y) in enumerate(dataloader):
        X, y = X.to(device, non_blocking=True), y.to(device, non_blocking=True)

        y_pred = model(X)
        loss = loss_fn(y_pred, y)
        loss_sum += loss.detach()
        acc_sum += _batch_correct(loss_fn, y_pred.detach(), y) / len(y_pred)

        optimizer.zero_grad()
        loss.backward()

        if fused_opt:
            optimizer.step(max_norm=MAX_GRAD_NORM)
        else:
            torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=MAX_GRAD_NORM)
            optimizer.step()
        lr_scheduler.step()

    n = len(dataloader)
    return loss_sum.item() / n, acc_sum.item() / n


def test_step(model: torch.nn.Module,
              dataloader: torch.utils.data.DataLoader,
              loss_fn: torch.nn.Module,
              device: torch.device) -> Tuple[float, float]:
    """One evaluation pass; returns (mean batch loss, mean batch accuracy)."""
    model.eval()
    loss_sum = torch.zeros((), dtype=torch.float32, device=device)
    acc_sum = torch.zeros((), dtype=torch.float32, device=device)
    with torch.inference_mode():
        for batch, (X, y) in enumerate(dataloader):
            X, y = X.to(device, non_blocking=True), y.to(device, non_blocking=True)
            logits = model(X)
            loss = loss_fn(logits, y)
            loss_sum += loss
            acc_sum += _batch_correct(loss_fn, logits, y) / len(logits)
    n = len(dataloader)
    return loss_sum.item() / n, acc_sum.item() / n


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
