"""engine.test_step and engine.train on the CUDA path (reference going_modular/going_modular/engine.py:81-130,
132-211) against the values the UNMODIFIED reference engine returned for the same model, loaders and recipe
(tests/golden/engine_tiny.npz, written by oracle/make_golden.py). Both precision modes; the fused loss / optimizer
and the stock torch ones the reference would pass."""
import numpy as np
import pytest
import torch

from common import TINY, engine_loaders, load_npz, notebook_recipe, sub

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def eng():
    return load_npz("engine_tiny.npz")


def build():
    from b200vit.models.vit import ViT
    m = ViT(**TINY)
    m.load_state_dict(sub(load_npz("vit_tiny_train.npz"), "sd0/"))
    return m


# tolerance on a loss value: north_star's 1 % for the bf16 path, 1e-4 in fp32 mode; accuracy may move by one
# sample of one batch when bf16 rounding flips a near-tie (batches of 8, 8, 5 -> at most 1/5 of a third)
TOL = {"bf16": (1e-2, 1.0 / 15 + 1e-9), "fp32": (1e-4, 1e-9)}


@pytest.mark.parametrize("mode", ["bf16", "fp32"])
@pytest.mark.parametrize("fused_loss", [True, False])
def test_test_step_matches_reference_engine(eng, mode, fused_loss):
    import b200vit
    from b200vit.going_modular import engine
    from b200vit.nn import CrossEntropyLoss
    _, test = engine_loaders(eng)
    m = build().to(dev())
    m.train()                                    # test_step must switch to eval itself (engine.py:104)
    loss_fn = CrossEntropyLoss() if fused_loss else torch.nn.CrossEntropyLoss()
    with b200vit.precision(mode):
        loss, acc = engine.test_step(model=m, dataloader=test, loss_fn=loss_fn, device=dev())
    assert isinstance(loss, float) and isinstance(acc, float)
    assert not m.training
    ltol, atol = TOL[mode]
    assert abs(loss - float(eng["test_step_loss"])) < ltol * float(eng["test_step_loss"])
    assert abs(acc - float(eng["test_step_acc"])) <= atol
    for p in m.parameters():
        assert p.grad is None                    # inference_mode: nothing recorded, nothing accumulated


@pytest.mark.parametrize("mode", ["bf16", "fp32"])
def test_train_matches_reference_engine(eng, mode, capsys):
    """3 epochs of engine.train: model moved to the device by train() itself (engine.py:180), per-epoch print line,
    results dict with the reference's keys and values."""
    import b200vit
    from b200vit.going_modular import engine
    from b200vit.nn import CrossEntropyLoss
    train, test = engine_loaders(eng)
    m = build()                                  # still on the CPU: train() does model.to(device)
    opt, sched = notebook_recipe(m, int(eng["total_steps"]), fused=True)
    with b200vit.precision(mode):
        res = engine.train(model=m, train_dataloader=train, test_dataloader=test, optimizer=opt,
                           loss_fn=CrossEntropyLoss(), lr_scheduler=sched, epochs=int(eng["epochs"]), device=dev())
    assert next(m.parameters()).is_cuda
    assert list(res) == ["train_loss", "train_acc", "test_loss", "test_acc"]
    ltol, atol = TOL[mode]
    for k in ("train_loss", "test_loss"):
        assert len(res[k]) == int(eng["epochs"])
        np.testing.assert_allclose(res[k], eng["train/" + k], rtol=ltol)
    np.testing.assert_allclose(res["test_acc"], eng["train/test_acc"], atol=atol)
    np.testing.assert_allclose(res["train_acc"], eng["train/train_acc"], atol=1.0 / 24 + 1e-9 if mode == "bf16" else 1e-9)
    out = capsys.readouterr().out
    assert f"Epoch: 1 | train_loss: {res['train_loss'][0]:.4f} | train_acc: {res['train_acc'][0]:.4f} | " \
           f"test_loss: {res['test_loss'][0]:.4f} | test_acc: {res['test_acc'][0]:.4f}" in out


def test_train_with_stock_torch_optimizer_and_loss(eng):
    """The reference passes torch.optim.Adam and nn.CrossEntropyLoss (notebook :2818, :3019): the same call works
    here and takes the clip_grad_norm_ + optimizer.step() route of engine.py:63-66."""
    from b200vit.going_modular import engine
    train, test = engine_loaders(eng)
    m = build()
    opt, sched = notebook_recipe(m, int(eng["total_steps"]), fused=False)
    res = engine.train(model=m, train_dataloader=train, test_dataloader=test, optimizer=opt,
                       loss_fn=torch.nn.CrossEntropyLoss(), lr_scheduler=sched, epochs=int(eng["epochs"]), device=dev())
    for k in ("train_loss", "test_loss"):
        np.testing.assert_allclose(res[k], eng["train/" + k], rtol=1e-2)
