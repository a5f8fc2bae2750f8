"""CPU tests: the oracle restatement against golden vectors produced by the unmodified reference
(oracle/make_golden.py), and the drop-in boundary (constructors, state_dict keys, RNG order)."""
import numpy as np
import pytest
import torch

from common import TINY, load_npz, rel_err, sub
from oracle import vit_oracle as O


@pytest.fixture(scope="module")
def tiny():
    return load_npz("vit_tiny_train.npz")


def tiny_cfg(**kw):
    d = dict(TINY)
    d.update(kw)
    return O.ViTConfig(**d)


def test_oracle_forward_matches_reference(tiny):
    sd = sub(tiny, "sd0/")
    X = torch.from_numpy(tiny["X"])
    logits = O.vit_forward(X, sd, tiny_cfg())
    # fp32 CPU restatement vs fp32 CPU reference: summation-order noise only
    assert rel_err(logits, torch.from_numpy(tiny["logits_eval"])) < 1e-5


def test_oracle_loss_and_grads_match_reference(tiny):
    sd = {k: v.clone().requires_grad_(True) for k, v in sub(tiny, "sd0/").items()}
    X = torch.from_numpy(tiny["X"])
    y = torch.from_numpy(tiny["y"])
    loss = O.cross_entropy(O.vit_forward(X, sd, tiny_cfg()), y)
    assert abs(loss.item() - float(tiny["loss"])) < 1e-5
    loss.backward()
    ref = sub(tiny, "grad/")
    assert set(ref) == set(sd)
    for k, g in ref.items():
        assert rel_err(sd[k].grad, g) < 2e-4, k


def test_oracle_train_steps_match_reference_engine(tiny):
    """engine.train_step of the reference (Adam coupled-L2 + clip 1.0 + warm-up/decay schedule) vs OracleTrainer."""
    tr = O.OracleTrainer(sub(tiny, "sd0/"), tiny_cfg(), total_steps=int(tiny["total_steps"]))
    losses, accs = [], []
    for X, y in zip(tiny["batches_X"], tiny["batches_y"]):
        l, a = tr.step(torch.from_numpy(X), torch.from_numpy(y))
        losses.append(l)
        accs.append(a)
    np.testing.assert_allclose(losses, tiny["traj_loss"], rtol=2e-4)
    np.testing.assert_allclose(accs, tiny["traj_acc"], atol=1e-12)
    final = sub(tiny, "sd_final/")
    for k, v in final.items():
        # key-bias entries of in_proj_bias have a mathematically zero gradient (softmax shift invariance): Adam
        # normalises their round-off noise into O(lr) steps, so allow a small absolute slack there
        if k.endswith("in_proj_bias"):
            assert (tr.sd[k].detach() - v).abs().max().item() < 5e-3, k    # <= sum of the learning rates used
        else:
            assert rel_err(tr.sd[k], v) < 1e-4, k


def test_oracle_test_step_and_train_match_reference_engine(tiny):
    """engine.test_step (engine.py:81-130) and engine.train (engine.py:132-211) of the unmodified reference, recorded
    in tests/golden/engine_tiny.npz, vs the oracle's restatements."""
    from common import engine_loaders
    eng = load_npz("engine_tiny.npz")
    train, test = engine_loaders(eng)
    assert [len(y) for _, y in train] == [8, 8, 8, 6] and [len(y) for _, y in test] == [8, 8, 5]
    sd = sub(tiny, "sd0/")
    l, a = O.evaluate(sd, tiny_cfg(), test)
    assert abs(l - float(eng["test_step_loss"])) < 1e-5 and a == pytest.approx(float(eng["test_step_acc"]), abs=1e-12)
    res = O.train(sd, tiny_cfg(), train, test, int(eng["epochs"]), int(eng["total_steps"]))
    for k in ("train_loss", "test_loss"):
        np.testing.assert_allclose(res[k], eng["train/" + k], rtol=2e-4)
    for k in ("train_acc", "test_acc"):
        np.testing.assert_allclose(res[k], eng["train/" + k], atol=1e-12)


def test_lr_schedule_matches_torch_sequential_lr():
    lin = torch.nn.Linear(2, 2)
    opt = torch.optim.Adam(lin.parameters(), lr=1e-3)
    total = 40
    warmup = int(0.05 * total)
    warm = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1e-6, end_factor=1.0, total_iters=warmup)
    dec = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1.0, end_factor=0.0, total_iters=total - warmup)
    sched = torch.optim.lr_scheduler.SequentialLR(opt, schedulers=[warm, dec], milestones=[warmup])
    for step in range(total):
        assert opt.param_groups[0]["lr"] == pytest.approx(O.lr_at(step, total), rel=1e-9, abs=1e-15), step
        opt.step()
        sched.step()


def test_headless_oracle_matches_reference(tiny):
    feats = load_npz("headless_tiny.npz")["feats"]
    sd = {k: v for k, v in sub(tiny, "sd0/").items() if not k.startswith("classifier")}
    out = O.vit_forward(torch.from_numpy(tiny["X"]), sd, tiny_cfg(num_classes=None))
    assert out.shape == feats.shape
    assert rel_err(out, torch.from_numpy(feats)) < 1e-5


def test_decay_split_counts():
    from b200vit.models.vit import ViT
    m = ViT(num_transformer_layer=12, num_classes=3)
    decay, no_decay = O.decay_split(m.named_parameters())
    assert (len(decay), len(no_decay)) == (52, 100)     # counted on the reference
    assert "patch_embedding_block.class_token" in decay and "patch_embedding_block.position_embedding" in decay


# ---- drop-in boundary ---------------------------------------------------------------------------------
def test_constructor_reproduces_reference_weights_and_kat():
    """torch.manual_seed(42) + b200vit's ViT(num_classes=3) must give the reference's initial weights (checked
    through per-tensor checksums recorded from the reference) and, pushed through the oracle, the notebook's
    known-answer logits (train mode) and the reference's eval-mode logits."""
    from b200vit.models.vit import ViT
    kat = load_npz("vit_b16_kat.npz")
    torch.manual_seed(42)
    x = torch.randn(1, 3, 224, 224)
    m = ViT(num_classes=3)
    sd = m.state_dict()
    chk = {k[4:]: v for k, v in kat.items() if k.startswith("chk/")}
    assert list(sd.keys()) == list(chk.keys()) and len(sd) == 152
    assert sum(v.numel() for v in sd.values()) == 85_800_963       # notebook :2727
    for k, v in sd.items():
        assert v.dtype == torch.float32
        np.testing.assert_allclose([float(v.double().sum()), float(v.double().abs().sum())], chk[k], rtol=1e-12, err_msg=k)
    cfg = O.ViTConfig(num_classes=3)
    out_train = O.vit_forward(x, sd, cfg, training=True)            # consumes the CPU generator like nn.Dropout
    np.testing.assert_allclose(out_train.numpy(), kat["logits_train"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(out_train.numpy().round(4), kat["notebook"], atol=1.01e-4)
    out_eval = O.vit_forward(x, sd, cfg, training=False)
    np.testing.assert_allclose(out_eval.numpy(), kat["logits_eval"], rtol=1e-4, atol=1e-5)


def test_constructor_signatures_match_reference():
    import inspect
    from b200vit.models import vit, vit_no_classifier
    from b200vit.going_modular import engine

    def params(fn):
        return [(p.name, p.default) for p in inspect.signature(fn).parameters.values() if p.name != "self"]

    E = inspect.Parameter.empty
    assert params(vit.PatchEmbedding.__init__) == [("image_size", E), ("color_channels", 3), ("patch_size", 16),
                                                    ("embedding_dropout", 0.1), ("embedding_dim", 768)]
    assert params(vit.MultiHeadSelfAttentionBlock.__init__) == [("embedding_dim", 768), ("num_heads", 12), ("attn_dropout", 0)]
    assert params(vit.MLPBlock.__init__) == [("embedding_dim", 768), ("mlp_size", 3072), ("dropout", 0.1)]
    assert params(vit.TransformerEncoderBlock.__init__) == [("embedding_dim", 768), ("num_heads", 12), ("attn_dropout", 0),
                                                             ("mlp_size", 3072), ("mlp_dropout", 0.1)]
    vit_args = [("image_size", 224), ("patch_size", 16), ("num_transformer_layer", 12), ("num_heads", 12),
                ("embedding_dim", 768), ("mlp_size", 3072), ("attn_dropout", 0), ("mlp_dropout", 0.1),
                ("embedding_dropout", 0.1)]
    assert params(vit.ViT.__init__) == vit_args + [("num_classes", 1000)]
    assert params(vit_no_classifier.ViT.__init__) == vit_args
    # the reference's positional parameters, in order; anything more must be keyword-only with a default (an extension
    # that a reference caller never sees): accumulation_steps
    sig = inspect.signature(engine.train_step).parameters.values()
    assert [p.name for p in sig if p.kind != inspect.Parameter.KEYWORD_ONLY] == \
        ["model", "dataloader", "loss_fn", "optimizer", "lr_scheduler", "device"]
    assert all(p.default is not E for p in sig if p.kind == inspect.Parameter.KEYWORD_ONLY)
    assert [n for n, _ in params(engine.test_step)] == ["model", "dataloader", "loss_fn", "device"]
    assert [n for n, _ in params(engine.train)] == ["model", "train_dataloader", "test_dataloader", "optimizer",
                                                    "loss_fn", "lr_scheduler", "epochs", "device"]


def test_no_cpu_fallback():
    """The product path must fail loudly on CPU tensors instead of computing something."""
    from b200vit.models.vit import ViT
    m = ViT(**TINY)
    with pytest.raises(RuntimeError, match="CUDA|libb200vit"):
        m(torch.rand(1, 3, 64, 64))
