"""Fused loss module usable wherever the reference passes `loss_fn` (nn.CrossEntropyLoss() in the notebook,
called at going_modular/engine.py:53,120)."""
from __future__ import annotations

import torch
from torch import nn

from . import functional as F_


class CrossEntropyLoss(nn.Module):
    """Mean softmax cross-entropy over [batch, classes] logits and int64 class labels, fused with its
    gradient and with the correct-prediction count the engine needs for accuracy (engine.py:73-74).

    After a call, `last_correct` holds a device int32 tensor with the number of argmax hits and
    `last_batch` the batch size; the engine reads them instead of launching softmax/argmax/compare."""

    def __init__(self):
        super().__init__()
        self._stats = {}

    @property
    def last_correct(self):
        return self._stats.get("correct")

    @property
    def last_batch(self):
        return self._stats.get("batch")

    def forward(self, input: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        return F_.CrossEntropyFn.apply(input, target, self._stats)
