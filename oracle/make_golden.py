"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

    python oracle/make_golden.py            # needs /root/reference (read-only mount)

The reference is Python and cannot travel to the GPU box, so its outputs are frozen here as small fixtures:

  vit_tiny_train.npz   tiny ViT (image 64, patch 16, D=128, 2 heads, 2 layers, mlp 256, 10 classes):
                       initial state_dict (seed 42), one batch (X, y), eval-mode logits, loss, every parameter
                       gradient of one backward pass, and the loss/accuracy trajectory + final weights of
                       6 engine.train_step iterations (reference engine.py, Adam + clip + SequentialLR recipe of
                       the notebook, dropout 0).
  vit_b16_kat.npz      ViT-B/16 3-class, seed 42: the notebook's known-answer logits (train mode, dropout on,
                       CPU generator) and the eval-mode logits for the same input; input is regenerated from the
                       seed, weights are NOT stored (343 MB) -- they are re-created by seeding a constructor with
                       the same RNG consumption (tests check b200vit's constructor reproduces them).
  headless_tiny.npz    models/vit_no_classifier.py forward on the tiny config (weights = vit_tiny_train sd0 minus
                       the classifier: asserted identical here).
  engine_tiny.npz      the reference's own engine.test_step (engine.py:81-130) on the tiny model at its initial weights
                       over a ragged 3-batch loader, and engine.train (engine.py:132-211) for 3 epochs (4 train
                       batches, the same test loader): the four metric lists it returns.   (`--only engine` writes
                       just this file.)
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"

TINY = dict(image_size=64, patch_size=16, num_transformer_layer=2, num_heads=2, embedding_dim=128, mlp_size=256,
            attn_dropout=0, mlp_dropout=0.0, embedding_dropout=0.0, num_classes=10)


def notebook_recipe(model, total_steps):
    """Optimizer / scheduler / loss exactly as in 08_pytorch_paper_replicating_self.ipynb :2791-2823, :2941-2960,
    :3019."""
    decay, no_decay = [], []
    for name, param in model.named_parameters():
        if not param.requires_grad:
            continue
        if param.ndim == 1 or name.endswith(".bias"):
            no_decay.append(param)
        else:
            decay.append(param)
    optimizer = torch.optim.Adam([{"params": decay, "weight_decay": 0.03}, {"params": no_decay, "weight_decay": 0.0}],
                                 lr=1e-3, betas=(0.9, 0.999))
    warmup = int(0.05 * total_steps)
    warm = torch.optim.lr_scheduler.LinearLR(optimizer, start_factor=1e-6, end_factor=1.0, total_iters=warmup)
    dec = torch.optim.lr_scheduler.LinearLR(optimizer, start_factor=1.0, end_factor=0.0,
                                            total_iters=total_steps - warmup)
    sched = torch.optim.lr_scheduler.SequentialLR(optimizer, schedulers=[warm, dec], milestones=[warmup])
    return optimizer, sched, torch.nn.CrossEntropyLoss()


def engine_fixture(ViT, engine):
    """engine.test_step / engine.train of the unmodified reference on the tiny model (seed 42 weights = sd0 of
    vit_tiny_train.npz)."""
    torch.manual_seed(42)
    model = ViT(**TINY)
    g = torch.Generator().manual_seed(11)
    sizes_train, sizes_test = [8, 8, 8, 6], [8, 8, 5]          # ragged last batches: mean of per-batch means
    train = [(torch.rand(n, 3, 64, 64, generator=g), torch.randint(0, 10, (n,), generator=g)) for n in sizes_train]
    test = [(torch.rand(n, 3, 64, 64, generator=g), torch.randint(0, 10, (n,), generator=g)) for n in sizes_test]
    loss_fn = torch.nn.CrossEntropyLoss()
    t_loss, t_acc = engine.test_step(model=model, dataloader=test, loss_fn=loss_fn, device="cpu")
    epochs = 3
    total_steps = 40                    # schedule horizon: 2 warm-up steps, then the decay (12 steps are run)
    optimizer, sched, loss_fn = notebook_recipe(model, total_steps)
    res = engine.train(model=model, train_dataloader=train, test_dataloader=test, optimizer=optimizer, loss_fn=loss_fn,
                       lr_scheduler=sched, epochs=epochs, device="cpu")
    out = dict(test_step_loss=np.float64(t_loss), test_step_acc=np.float64(t_acc), epochs=np.int64(epochs),
               total_steps=np.int64(total_steps))
    for k, v in res.items():
        out["train/" + k] = np.array(v, dtype=np.float64)
    for i, (X, y) in enumerate(train):
        out[f"train_X{i}"], out[f"train_y{i}"] = X.numpy(), y.numpy()
    for i, (X, y) in enumerate(test):
        out[f"test_X{i}"], out[f"test_y{i}"] = X.numpy(), y.numpy()
    np.savez_compressed(OUT / "engine_tiny.npz", **out)
    print("engine fixture: test_step", t_loss, t_acc, "train", res)


def main():
    if not REF.exists():
        sys.exit("reference not mounted at /root/reference")
    sys.path.insert(0, str(REF))
    sys.path.insert(0, str(REF / "going_modular"))
    from models.vit import ViT  # noqa
    from models import vit_no_classifier  # noqa
    from going_modular import engine  # noqa

    OUT.mkdir(parents=True, exist_ok=True)
    torch.set_num_threads(8)
    if "--only" in sys.argv and sys.argv[sys.argv.index("--only") + 1] == "engine":
        engine_fixture(ViT, engine)
        return

    # ---- tiny training fixture -------------------------------------------------------------------
    torch.manual_seed(42)
    model = ViT(**TINY)
    sd0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
    g = torch.Generator().manual_seed(0)
    X = torch.rand(8, 3, 64, 64, generator=g)
    y = torch.randint(0, 10, (8,), generator=g)
    model.eval()
    with torch.no_grad():
        logits_eval = model(X)
    model.train()  # dropout is 0 in TINY, so train == eval numerically
    logits = model(X)
    loss = torch.nn.functional.cross_entropy(logits, y)
    loss.backward()
    grads = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
    model.zero_grad()

    steps = 6
    total_steps = 40
    optimizer, sched, loss_fn = notebook_recipe(model, total_steps)
    batches = []
    for _ in range(steps):
        batches.append((torch.rand(8, 3, 64, 64, generator=g), torch.randint(0, 10, (8,), generator=g)))
    traj_loss, traj_acc = [], []
    for b in batches:  # one train_step call per batch so that every per-step loss is recorded
        l, a = engine.train_step(model=model, dataloader=[b], loss_fn=loss_fn, optimizer=optimizer,
                                 lr_scheduler=sched, device="cpu")
        traj_loss.append(l)
        traj_acc.append(a)
    out = {f"sd0/{k}": v.numpy() for k, v in sd0.items()}
    out.update({f"grad/{k}": v.numpy() for k, v in grads.items()})
    out.update({f"sd_final/{k}": v.detach().numpy() for k, v in model.state_dict().items()})
    out.update(X=X.numpy(), y=y.numpy(), logits_eval=logits_eval.numpy(), loss=np.float32(loss.item()),
               traj_loss=np.array(traj_loss, dtype=np.float64), traj_acc=np.array(traj_acc, dtype=np.float64),
               batches_X=np.stack([b[0].numpy() for b in batches]), batches_y=np.stack([b[1].numpy() for b in batches]),
               total_steps=np.int64(total_steps))
    np.savez_compressed(OUT / "vit_tiny_train.npz", **out)

    # ---- headless tiny ---------------------------------------------------------------------------
    torch.manual_seed(42)
    cfg = {k: v for k, v in TINY.items() if k != "num_classes"}
    hm = vit_no_classifier.ViT(**cfg).eval()
    with torch.no_grad():
        feats = hm(X)
    for k, v in hm.state_dict().items():   # same seed + same RNG order => same weights as the classifier model
        assert torch.equal(v, sd0[k]), k
    np.savez_compressed(OUT / "headless_tiny.npz", feats=feats.numpy())

    # ---- ViT-B/16 known-answer test (notebook cell 78, :2605-2631) --------------------------------
    torch.manual_seed(42)
    xk = torch.randn(1, 3, 224, 224)
    vit = ViT(num_classes=3)
    out_train = vit(xk).detach()          # fresh modules are in train mode: dropout active, as in the notebook
    vit.eval()
    with torch.no_grad():
        out_eval = vit(xk)
    sd = vit.state_dict()
    checks = {k: np.array([float(v.double().sum()), float(v.double().abs().sum())]) for k, v in sd.items()}
    np.savez_compressed(OUT / "vit_b16_kat.npz", logits_train=out_train.numpy(), logits_eval=out_eval.numpy(),
                        notebook=np.array([[0.6136, -0.9092, 0.5918]], dtype=np.float32),
                        **{f"chk/{k}": v for k, v in checks.items()})
    print("train-mode logits", out_train, "eval", out_eval)
    engine_fixture(ViT, engine)
    print("written to", OUT)


if __name__ == "__main__":
    main()
