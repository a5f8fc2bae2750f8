"""CUDA-graph replay of the eval-mode forward (b200vit.inference.GraphedForward) against the eager forward."""
import pytest
import torch

from common import TINY, rel_err

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


@pytest.mark.parametrize("headless", [False, True])
def test_graphed_forward_matches_eager(headless):
    from b200vit.inference import GraphedForward
    from b200vit.models import vit, vit_no_classifier
    torch.manual_seed(3)
    kw = dict(TINY)
    if headless:
        kw.pop("num_classes")
        model = vit_no_classifier.ViT(**kw)
    else:
        model = vit.ViT(**kw)
    model = model.to(_dev()).eval()
    fwd = GraphedForward(model)
    g = torch.Generator().manual_seed(5)
    for B in (1, 3, 1, 3):                                   # second visit of a shape replays the recorded graph
        x = torch.rand(B, 3, TINY["image_size"], TINY["image_size"], generator=g).to(_dev())
        with torch.inference_mode():
            ref = model(x).clone()
        got = fwd(x).clone()
        assert got.shape == ref.shape
        assert rel_err(got, ref) < 1e-5
    assert len(fwd._graphs) == 2

    # in-place parameter updates must be visible to the replay (bf16 shadows are re-cast before it)
    with torch.no_grad():
        for p in model.parameters():
            p.mul_(1.05)
    x = torch.rand(3, 3, TINY["image_size"], TINY["image_size"], generator=g).to(_dev())
    with torch.inference_mode():
        ref = model(x).clone()
    got = fwd(x).clone()
    assert rel_err(got, ref) < 1e-5
    assert len(fwd._graphs) == 2


def test_graphed_forward_rejects_training_mode():
    from b200vit.inference import GraphedForward
    from b200vit.models import vit
    model = vit.ViT(**TINY).to(_dev()).train()
    fwd = GraphedForward(model)
    with pytest.raises(RuntimeError):
        fwd(torch.rand(1, 3, TINY["image_size"], TINY["image_size"], device=_dev()))
