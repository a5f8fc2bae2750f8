"""Shared helpers for the test-suite."""
from pathlib import Path

import numpy as np
import torch

GOLDEN = Path(__file__).resolve().parent / "golden"

TINY = dict(image_size=64, patch_size=16, num_transformer_layer=2, num_heads=2, embedding_dim=128, mlp_size=256,
            attn_dropout=0, mlp_dropout=0.0, embedding_dropout=0.0, num_classes=10)


def load_npz(name):
    z = np.load(GOLDEN / name)
    return {k: z[k] for k in z.files}


def sub(d, prefix):
    return {k[len(prefix):]: torch.from_numpy(np.array(v)) for k, v in d.items() if k.startswith(prefix)}


def rel_err(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def notebook_recipe(model, total_steps, fused=False, lr=1e-3):
    """Optimizer / scheduler of the reference notebook (:2791-2823, :2941-2960). fused=True swaps in FusedAdam."""
    decay, no_decay = [], []
    for name, param in model.named_parameters():
        if param.ndim == 1 or name.endswith(".bias"):
            no_decay.append(param)
        else:
            decay.append(param)
    groups = [{"params": decay, "weight_decay": 0.03}, {"params": no_decay, "weight_decay": 0.0}]
    if fused:
        from b200vit.optim import FusedAdam
        opt = FusedAdam(groups, lr=lr, betas=(0.9, 0.999))
    else:
        opt = torch.optim.Adam(groups, lr=lr, betas=(0.9, 0.999))
    warmup = int(0.05 * total_steps)
    warm = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1e-6, end_factor=1.0, total_iters=warmup)
    dec = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1.0, end_factor=0.0, total_iters=total_steps - warmup)
    sched = torch.optim.lr_scheduler.SequentialLR(opt, schedulers=[warm, dec], milestones=[warmup])
    return opt, sched


def engine_loaders(eng):
    """(train batches, test batches) of tests/golden/engine_tiny.npz as lists of (X, y) CPU tensors."""
    def grab(prefix):
        out, i = [], 0
        while f"{prefix}_X{i}" in eng:
            out.append((torch.from_numpy(eng[f"{prefix}_X{i}"]), torch.from_numpy(eng[f"{prefix}_y{i}"])))
            i += 1
        return out
    return grab("train"), grab("test")
