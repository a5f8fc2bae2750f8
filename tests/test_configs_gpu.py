"""Parity cases for the other BASELINE.json configs (they are test cases, not bench lines):
config 2 (1000-class head), config 3 (ViT-L/16 @384: 577 tokens, d=1024, 16 heads -> long-sequence attention path,
LayerNorm D=1024), config 4 (headless forward-only sweep over batch sizes). The 200-step loss curves are in test_loss_curve_gpu.py."""
import numpy as np
import pytest
import torch

from common import TINY, load_npz, rel_err, sub
from oracle import vit_oracle as O

pytestmark = pytest.mark.gpu
BF16_TOL = 2e-2


def dev():
    return torch.device("cuda:0")


def test_config2_vit_b16_1000_classes_logits_and_head_grads():
    from b200vit.models.vit import ViT
    torch.manual_seed(1)
    kw = dict(num_transformer_layer=2, mlp_dropout=0.0, embedding_dropout=0.0, num_classes=1000)
    m = ViT(**kw)
    sd = {k: v.clone().requires_grad_(True) for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(2)
    X = torch.rand(3, 3, 224, 224, generator=g)
    y = torch.randint(0, 1000, (3,), generator=g)
    ref = O.vit_forward(X, sd, O.ViTConfig(**kw))
    O.cross_entropy(ref, y).backward()
    m = m.to(dev()).train()
    out = m(X.to(dev()))
    assert out.shape == (3, 1000)
    assert rel_err(out, ref) < BF16_TOL
    torch.nn.functional.cross_entropy(out, y.to(dev())).backward()
    for k in ("classifier.0.weight", "classifier.0.bias", "layer_norm.weight", "patch_embedding_block.position_embedding"):
        p = dict(m.named_parameters())[k]
        assert rel_err(p.grad, sd[k].grad) < 5e-2, k


def test_config3_vit_l16_384_shape_two_layers():
    """ViT-L/16 geometry (577 tokens, d=1024, 16 heads, mlp 4096) with 2 layers: exercises the T > 256 attention
    kernels, LayerNorm at D=1024 and the K=1024/4096 GEMMs, forward and backward."""
    from b200vit.models.vit import ViT
    torch.manual_seed(3)
    kw = dict(image_size=384, num_transformer_layer=2, num_heads=16, embedding_dim=1024, mlp_size=4096,
              mlp_dropout=0.0, embedding_dropout=0.0, num_classes=10)
    m = ViT(**kw)
    sd = {k: v.clone().requires_grad_(True) for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(4)
    X = torch.rand(2, 3, 384, 384, generator=g)
    y = torch.randint(0, 10, (2,), generator=g)
    ref = O.vit_forward(X, sd, O.ViTConfig(**kw))
    O.cross_entropy(ref, y).backward()
    m = m.to(dev()).train()
    out = m(X.to(dev()))
    assert rel_err(out, ref) < BF16_TOL
    torch.nn.functional.cross_entropy(out, y.to(dev())).backward()
    bad = {}
    for k, p in m.named_parameters():
        if k.endswith("in_proj_bias"):
            continue
        e = rel_err(p.grad, sd[k].grad)
        if e > 6e-2:
            bad[k] = e
    assert not bad, bad


@pytest.mark.parametrize("batch", [1, 2, 5, 16])
def test_config4_headless_forward_sweep(batch):
    from b200vit.models import vit_no_classifier
    tiny = load_npz("vit_tiny_train.npz")
    cfg = {k: v for k, v in TINY.items() if k != "num_classes"}
    m = vit_no_classifier.ViT(**cfg)
    sd = {k: v for k, v in sub(tiny, "sd0/").items() if not k.startswith("classifier")}
    m.load_state_dict(sd)
    m = m.to(dev()).eval()
    g = torch.Generator().manual_seed(100 + batch)
    X = torch.rand(batch, 3, 64, 64, generator=g)
    ref = O.vit_forward(X, sd, O.ViTConfig(**cfg, num_classes=None))
    with torch.inference_mode():
        out = m(X.to(dev()))
    assert out.shape == (batch, 17, 128)
    assert rel_err(out, ref) < BF16_TOL


def test_headless_vit_b16_forward_b1():
    from b200vit.models import vit_no_classifier
    torch.manual_seed(5)
    m = vit_no_classifier.ViT(mlp_dropout=0.0, embedding_dropout=0.0)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    X = torch.rand(1, 3, 224, 224, generator=torch.Generator().manual_seed(6))
    ref = O.vit_forward(X, sd, O.ViTConfig(mlp_dropout=0.0, embedding_dropout=0.0, num_classes=None))
    with torch.inference_mode():
        out = m.to(dev()).eval()(X.to(dev()))
    assert out.shape == (1, 197, 768)
    assert rel_err(out, ref) < BF16_TOL
