"""Edge cases of the boundary (reference models/vit.py:25, 45-67, 224-236; engine.py:45-79): a single image, two tokens
per image, patch sizes other than 16, a wrong image size, ragged last batches, a very large batch of tiny images."""
import pytest
import torch

from common import rel_err
from oracle import vit_oracle as O

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


def _check(cfg, batch, seed=0, train=True):
    from b200vit.models.vit import ViT
    torch.manual_seed(seed)
    m = ViT(**cfg)
    sd = {k: v.clone().requires_grad_(True) for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(seed + 1)
    X = torch.rand(batch, 3, cfg["image_size"], cfg["image_size"], generator=g)
    y = torch.randint(0, cfg["num_classes"], (batch,), generator=g)
    ref = O.vit_forward(X, sd, O.ViTConfig(**cfg))
    m = m.to(dev()).train()
    out = m(X.to(dev()))
    assert out.shape == ref.shape
    assert rel_err(out, ref) < 2e-2
    if train:
        O.cross_entropy(ref, y).backward()
        torch.nn.functional.cross_entropy(out, y.to(dev())).backward()
        bad = {}
        for k, p in m.named_parameters():
            if k.endswith("in_proj_bias"):
                continue
            e = rel_err(p.grad, sd[k].grad)
            if e > 6e-2:
                bad[k] = e
        assert not bad, bad


BASE = dict(num_transformer_layer=2, num_heads=2, embedding_dim=128, mlp_size=256, attn_dropout=0, mlp_dropout=0.0,
            embedding_dropout=0.0, num_classes=5)


def test_single_image():
    _check(dict(BASE, image_size=64, patch_size=16), batch=1)


def test_two_tokens_per_image():
    _check(dict(BASE, image_size=16, patch_size=16), batch=3)          # one patch + the class token


def test_patch_size_8_and_32():
    _check(dict(BASE, image_size=32, patch_size=8), batch=2)           # 17 tokens of 192 values
    _check(dict(BASE, image_size=64, patch_size=32), batch=2)          # 5 tokens of 3072 values


def test_many_tiny_images():
    _check(dict(BASE, image_size=32, patch_size=16), batch=1000, train=False)     # 5 000 token rows of 5 tokens each


def test_wrong_image_size_is_rejected():
    """The reference fails at `+ position_embedding` when H x W does not match the constructor (vit.py:63); here the
    mismatch is reported before any kernel runs."""
    from b200vit.models.vit import ViT
    m = ViT(**dict(BASE, image_size=64, patch_size=16)).to(dev())
    with pytest.raises(RuntimeError):
        m(torch.rand(2, 3, 32, 32, device=dev()))
    with pytest.raises(AssertionError):
        ViT(**dict(BASE, image_size=60, patch_size=16))                 # reference assert, vit.py:25


def test_ragged_last_batch_through_the_engine():
    """engine.train_step averages per-batch means; a short last batch is weighted like a full one (engine.py:77-78)."""
    from b200vit.going_modular import engine
    from b200vit.models.vit import ViT
    from b200vit.nn import CrossEntropyLoss
    from b200vit.optim import FusedAdam
    cfg = dict(BASE, image_size=32, patch_size=16)
    torch.manual_seed(3)
    m = ViT(**cfg).to(dev())
    g = torch.Generator().manual_seed(4)
    data = [(torch.rand(n, 3, 32, 32, generator=g), torch.randint(0, 5, (n,), generator=g)) for n in (8, 8, 3, 1)]
    opt = FusedAdam(m.parameters(), lr=0.0)                              # lr 0: the weights stay put, losses are comparable
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0)
    loss, acc = engine.train_step(m, data, CrossEntropyLoss(), opt, sched, dev())
    m.eval()
    with torch.no_grad():
        per = [torch.nn.functional.cross_entropy(m(x.to(dev())), y.to(dev())).item() for x, y in data]
    assert abs(loss - sum(per) / len(per)) < 2e-2 * abs(loss)
    assert 0.0 <= acc <= 1.0
