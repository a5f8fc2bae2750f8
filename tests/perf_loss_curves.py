"""Calibration run for the 200-step loss-curve gates (not a pytest file): b200vit (bf16 and fp32 mode) against the
oracle executed on the same GPU in fp32, several repetitions, raw and smoothed relative deviations.

    python tests/perf_loss_curves.py [tiny|vitb] [reps]
"""
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT / "pytorch-vit-paper-replication_b200"), str(ROOT), str(ROOT / "tests")):
    sys.path.insert(0, p)

import b200vit  # noqa: E402
from b200vit.going_modular import engine  # noqa: E402
from b200vit.models.vit import ViT  # noqa: E402
from b200vit.nn import CrossEntropyLoss  # noqa: E402
from common import TINY, notebook_recipe  # noqa: E402
from oracle import vit_oracle as O  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "tiny"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
lr = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-3
f64 = len(sys.argv) > 4 and sys.argv[4] == "f64"
dev = torch.device("cuda:0")
torch.backends.cuda.matmul.allow_tf32 = False
steps = 200
if which == "tiny":
    kw, B, img, C = dict(TINY), 16, 64, 10
else:
    kw, B, img, C = dict(num_classes=3, mlp_dropout=0.0, embedding_dropout=0.0), 32, 224, 3
torch.manual_seed(42)
sd0 = {k: v.clone() for k, v in ViT(**kw).state_dict().items()}
g = torch.Generator().manual_seed(77)
pool = [(torch.rand(B, 3, img, img, generator=g), torch.randint(0, C, (B,), generator=g)) for _ in range(8)]
batches = [pool[i % 8] for i in range(steps)]
dbatches = [(x.to(dev), y.to(dev)) for x, y in batches[:8]]

tr = O.OracleTrainer({k: v.to(dev) for k, v in sd0.items()}, O.ViTConfig(**kw), total_steps=steps, base_lr=lr)
ref = np.array([tr.step(*dbatches[i % 8])[0] for i in range(steps)])
print(f"{which} lr {lr}: oracle-on-GPU loss {ref[0]:.4f} -> {ref[-1]:.4f} (min {ref.min():.4f})", flush=True)
out = {"oracle": ref.tolist()}
k8 = np.ones(8) / 8
if f64:
    class T64(O.OracleTrainer):
        def __init__(self, sd, *a, **k):
            super().__init__(sd, *a, **k)
            self.sd = {n: v.detach().double().requires_grad_(True) for n, v in self.sd.items()}
            self.m = {n: torch.zeros_like(v) for n, v in self.sd.items()}
            self.v = {n: torch.zeros_like(v) for n, v in self.sd.items()}
    t64 = T64({k: v.to(dev) for k, v in sd0.items()}, O.ViTConfig(**kw), total_steps=steps, base_lr=lr)
    r64 = np.array([t64.step(dbatches[i % 8][0].double(), dbatches[i % 8][1])[0] for i in range(steps)])
    nf = np.abs(ref - r64) / np.abs(r64)
    print(f"oracle fp32 vs oracle fp64 (the reference's own round-off floor): raw max {nf.max():.4%} (step {nf.argmax()}), "
          f"first-100 max {nf[:100].max():.4%}", flush=True)
    out["oracle_f64"] = r64.tolist()
if len(sys.argv) > 4 and sys.argv[4] == "autocast":
    # yardstick: the same oracle arithmetic under torch's bf16 autocast (cuBLAS bf16 GEMMs, fp32 master weights)
    ta = O.OracleTrainer({k: v.to(dev) for k, v in sd0.items()}, O.ViTConfig(**kw), total_steps=steps, base_lr=lr)
    ac = []
    for i in range(steps):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            ac.append(ta.step(*dbatches[i % 8])[0])
    ac = np.array(ac)
    rel = np.abs(ac - ref) / np.abs(ref)
    sm = np.abs(np.convolve(ac, k8, mode="valid") - np.convolve(ref, k8, mode="valid")) / np.convolve(ref, k8, mode="valid")
    print(f"torch bf16 autocast of the oracle vs fp32 oracle: raw max {rel.max():.4%} (step {rel.argmax()}), "
          f"smoothed max {sm.max():.4%}, final loss {ac[-1]:.4f}", flush=True)
    out["autocast"] = ac.tolist()
for mode in ("bf16", "fp32"):
    for rep in range(reps if mode == "bf16" else 1):
        m = ViT(**kw)
        m.load_state_dict(sd0)
        m = m.to(dev)
        opt, sched = notebook_recipe(m, steps, fused=True, lr=lr)
        loss_fn = CrossEntropyLoss()
        with b200vit.precision(mode):
            ours = np.array([engine.train_step(m, [dbatches[i % 8]], loss_fn, opt, sched, dev)[0] for i in range(steps)])
        rel = np.abs(ours - ref) / np.abs(ref)
        sm = np.abs(np.convolve(ours, k8, mode="valid") - np.convolve(ref, k8, mode="valid")) / np.convolve(ref, k8, mode="valid")
        print(f"{mode} rep {rep}: raw max {rel.max():.4%} (step {rel.argmax()}), first-100 max {rel[:100].max():.4%}, "
              f"smoothed max {sm.max():.4%}, final loss {ours[-1]:.4f}", flush=True)
        out[f"{mode}_{rep}"] = ours.tolist()
if Path("gpurun_out").is_dir():
    json.dump(out, open(f"gpurun_out/r02_loss_curves_{which}_lr{lr:g}.json", "w"))
