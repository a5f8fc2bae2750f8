"""Checkpoint interoperability with torchvision's VisionTransformer (the pretrained model the reference notebooks
fine-tune, 08_pytorch_paper_replicating_self.ipynb:4039-4129): same weights -> same outputs."""
import pytest
import torch

from common import rel_err

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


def test_torchvision_vit_b16_weights_load_and_match():
    torchvision = pytest.importorskip("torchvision")
    from b200vit.models.vit import ViT
    from b200vit.models import vit_no_classifier
    from b200vit.utils import load_torchvision_vit
    torch.manual_seed(11)
    tv = torchvision.models.vit_b_16(weights=None).to(_dev()).eval()          # random weights: no network here
    with torch.no_grad():                                                     # torchvision zero-inits the head
        tv.heads.head.weight.normal_(0, 0.02)
        tv.heads.head.bias.normal_(0, 0.02)
    ours = load_torchvision_vit(ViT(num_classes=1000), tv.state_dict()).to(_dev()).eval()
    assert ours.layer_norm.eps == 1e-6 and ours.transformer_encoder[3].mlp_block.layer_norm.eps == 1e-6
    x = torch.rand(4, 3, 224, 224, generator=torch.Generator().manual_seed(2)).to(_dev())
    with torch.no_grad():
        ref = tv(x)
        got = ours(x)
    assert rel_err(got, ref) < 2e-2                                           # bf16-mode tolerance of BASELINE.json
    # headless: the encoder output equals torchvision's encoder (all tokens, after the final LayerNorm)
    feats = load_torchvision_vit(vit_no_classifier.ViT(), tv.state_dict()).to(_dev()).eval()
    with torch.no_grad():
        t = tv._process_input(x)
        t = torch.cat([tv.class_token.expand(x.shape[0], -1, -1), t], dim=1)
        ref_tokens = tv.encoder(t)
        got_tokens = feats(x)
    assert rel_err(got_tokens, ref_tokens) < 2e-2
    # a new 3-class head for fine-tuning keeps its own initialisation
    ft = load_torchvision_vit(ViT(num_classes=3), tv.state_dict(), load_head=False)
    assert ft.classifier[0].weight.shape == (3, 768)
