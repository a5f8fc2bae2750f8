"""Low-latency inference: the forward of an eval-mode b200vit model captured in a CUDA graph.

Small batches are launch-bound: a ViT-B/16 forward is ~150 kernel launches, so at batch <= 16 the host's enqueue
time (~1.8 ms), not the GPU, sets the latency. `GraphedForward` records the launches once per input shape and
replays them with one call (the reference's single-image path is
going_modular/predictions.py:56-74 -- `model.eval()`, `torch.inference_mode()`, one image -- and the headless
feature-extraction sweep of models/vit_no_classifier.py:217-226).

    fwd = GraphedForward(model)            # model: b200vit ViT (classifier or headless) on a CUDA device
    logits = fwd(images)                   # first call per shape captures; later calls replay

Semantics match `model.eval(); with torch.inference_mode(): model(x)`: same kernels, same order, same results.
The returned tensor is the graph's static output buffer -- it is overwritten by the next call with the same
input shape; `.clone()` it to keep it. Parameter updates made in place (optimizer steps, `load_state_dict`) are
picked up: the bf16 weight shadows are refreshed before the replay when a parameter's version changed. Moving the
model (`.to()`), which re-allocates parameter storage, drops the recorded graphs.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch

from . import functional as F_
from . import kernels as K


class _Entry:
    __slots__ = ("graph", "static_in", "static_out")

    def __init__(self, graph, static_in, static_out):
        self.graph, self.static_in, self.static_out = graph, static_in, static_out


class GraphedForward:
    def __init__(self, model: torch.nn.Module, max_shapes: int = 16):
        self.model = model
        self.max_shapes = max_shapes
        self._graphs: Dict[Tuple, _Entry] = {}
        self._params = list(model.parameters())
        if not self._params or self._params[0].device.type != "cuda":
            raise RuntimeError("GraphedForward needs a model on a CUDA device")
        self._ptrs = self._ptr_key()
        self._versions = self._version_key()

    def _ptr_key(self):
        return tuple(p.data_ptr() for p in self._params)

    def _version_key(self):
        return tuple(p._version for p in self._params)

    def _sync_weights(self) -> None:
        ptrs = self._ptr_key()
        if ptrs != self._ptrs:                      # storage moved: recorded pointers are stale
            self._graphs.clear()
            self._ptrs = ptrs
        vers = self._version_key()
        if vers != self._versions:                  # in-place update: re-cast the bf16 shadows (in place) eagerly
            for p in self._params:
                if F_.shadow_entry(p) is not None:
                    F_.bf16_shadow(p)
            F_.refresh_folds()                      # ... and the LayerNorm-folded weights of the inference path
            self._versions = vers

    def _capture(self, x: torch.Tensor) -> _Entry:
        if len(self._graphs) >= self.max_shapes:
            self._graphs.pop(next(iter(self._graphs)))
        static_in = x.clone()
        side = torch.cuda.Stream(device=x.device)
        side.wait_stream(torch.cuda.current_stream(x.device))
        with torch.cuda.stream(side), torch.inference_mode():
            for _ in range(2):                      # warm-up: creates the weight shadows, sizes the allocator
                self.model(static_in)
        torch.cuda.current_stream(x.device).wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.inference_mode(), torch.cuda.graph(graph):
            static_out = self.model(static_in)
        self._versions = self._version_key()
        return _Entry(graph, static_in, static_out)

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        if self.model.training:
            raise RuntimeError("GraphedForward replays an eval-mode forward: call model.eval() first")
        if x.device != self._params[0].device:
            raise RuntimeError(f"input on {x.device}, model on {self._params[0].device}")
        self._sync_weights()
        key = (tuple(x.shape), x.dtype)
        entry = self._graphs.get(key)
        if entry is None:
            entry = self._graphs[key] = self._capture(x)
        entry.static_in.copy_(x, non_blocking=True)
        entry.graph.replay()
        return entry.static_out


def predict(model: torch.nn.Module, images: torch.Tensor, device=None, graphed: "GraphedForward | None" = None):
    """Class probabilities and top-1 labels for a batch of images: the numerical part of the reference's
    `pred_and_plot_image` (going_modular/predictions.py:56-66 -- `model.to(device)`, `model.eval()`,
    `torch.inference_mode()`, forward, `torch.softmax(dim=1)`, `torch.argmax(dim=1)`; loading / resizing the picture and
    plotting stay with the caller).

    images: float [B, 3, H, W] (already transformed) or uint8 [B, 3, H, W] (scaled by 1/255 on the device); host or
    device tensors. Returns (probs float32 [B, C], labels int64 [B]) on the device. Pass a `GraphedForward` of the same
    model as `graphed` to replay a recorded CUDA graph (small-batch latency)."""
    if device is not None:
        model.to(device)
    model.eval()
    dev = next(model.parameters()).device
    x = images.to(dev, non_blocking=True)
    with torch.inference_mode():
        logits = graphed(x) if graphed is not None else model(x)
        if logits.dim() != 2:
            raise ValueError("predict() needs a classifier model (logits [batch, classes])")
        logits = logits.float().contiguous()
        bn, c = logits.shape
        probs = torch.empty_like(logits)
        labels = torch.empty(bn, dtype=torch.int64, device=dev)
        K.softmax_top1(logits, probs, labels, bn, c)
    return probs, labels
