"""Drive the UNMODIFIED reference (models/vit.py + going_modular/engine.py) for timing.  MEASUREMENT INFRASTRUCTURE ONLY.

Used by bench.py for the `--impl reference` arm, the `cpu_baseline` block and the same-box `gpu_eager_baseline` block.
Nothing of b200vit is on this path: the reference's own ViT, the reference's own `engine.train_step`, stock
torch.optim.Adam / LinearLR / SequentialLR / nn.CrossEntropyLoss exactly as the notebook builds them
(08_pytorch_paper_replicating_self.ipynb:2791-2823, 2941-2960, 3019).

The reference is looked up at /root/reference (build container) or baseline/_ref (GPU box; oracle/install_reference.py).
"""
from __future__ import annotations

import os
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
CANDIDATES = [Path("/root/reference"), ROOT / "baseline" / "_ref"]


def reference_root():
    for c in CANDIDATES:
        if (c / "models" / "vit.py").exists() and (c / "going_modular" / "going_modular" / "engine.py").exists():
            return c
    return None


def import_reference():
    """(ViT class, engine module) of the unmodified reference, or raise FileNotFoundError."""
    root = reference_root()
    if root is None:
        raise FileNotFoundError("reference not found at /root/reference nor baseline/_ref "
                                "(run oracle/install_reference.py in the build container)")
    for p in (str(root), str(root / "going_modular")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import importlib
    vit = importlib.import_module("models.vit")
    engine = importlib.import_module("going_modular.engine")
    assert Path(vit.__file__).resolve().is_relative_to(root.resolve()), vit.__file__      # not b200vit's namesake
    assert Path(engine.__file__).resolve().is_relative_to(root.resolve()), engine.__file__
    return vit.ViT, engine


def notebook_recipe(model, total_steps):
    decay, no_decay = [], []
    for name, param in model.named_parameters():
        if not param.requires_grad:
            continue
        (no_decay if (param.ndim == 1 or name.endswith(".bias")) else decay).append(param)
    opt = torch.optim.Adam([{"params": decay, "weight_decay": 0.03}, {"params": no_decay, "weight_decay": 0.0}],
                           lr=1e-3, betas=(0.9, 0.999))
    warm = max(1, int(0.05 * total_steps))
    sched = torch.optim.lr_scheduler.SequentialLR(opt, schedulers=[
        torch.optim.lr_scheduler.LinearLR(opt, start_factor=1e-6, end_factor=1.0, total_iters=warm),
        torch.optim.lr_scheduler.LinearLR(opt, start_factor=1.0, end_factor=0.0, total_iters=total_steps - warm)],
        milestones=[warm])
    return opt, sched, torch.nn.CrossEntropyLoss()


def time_reference_train_step(device: str, batch: int, steps: int, warmup: int, classes: int = 3, autocast: bool = False,
                              threads: int = 0):
    """images/sec of the reference's engine.train_step on `device` ('cpu' or 'cuda:0'): ViT-B/16 (reference defaults,
    dropout 0.1/0.1/0 as shipped), synthetic batch, notebook recipe. One engine.train_step call over `steps` batches is
    timed by wall clock (the reference syncs on `.item()` twice per batch, engine.py:54,74)."""
    ViT, engine = import_reference()
    if device == "cpu":
        torch.set_num_threads(threads or (os.cpu_count() or 1))
    else:
        torch.backends.cuda.matmul.allow_tf32 = False       # the reference's own precision: plain fp32
        torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(42)
    model = ViT(num_classes=classes)
    opt, sched, loss_fn = notebook_recipe(model, 1000)
    g = torch.Generator().manual_seed(0)
    X = torch.rand(batch, 3, 224, 224, generator=g)
    y = torch.randint(0, classes, (batch,), generator=g)

    def run(n):
        if autocast:
            with torch.autocast(device.split(":")[0], dtype=torch.bfloat16):
                return engine.train_step(model=model, dataloader=[(X, y)] * n, loss_fn=loss_fn, optimizer=opt,
                                         lr_scheduler=sched, device=device)
        return engine.train_step(model=model, dataloader=[(X, y)] * n, loss_fn=loss_fn, optimizer=opt,
                                 lr_scheduler=sched, device=device)

    model.to(device)
    if warmup > 0:
        run(warmup)
    if device != "cpu":
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    loss, _ = run(steps)
    if device != "cpu":
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    out = {"value": round(batch * steps / dt, 3), "s_per_step": round(dt / steps, 4), "loss": round(float(loss), 4),
           "reference_root": str(reference_root())}
    del model, opt
    return out
