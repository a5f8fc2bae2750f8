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


def test_predict_matches_oracle_probabilities_and_labels():
    """predict() = eval forward + softmax + argmax (reference going_modular/predictions.py:56-66), eager and through a
    recorded CUDA graph, against the CPU oracle's logits pushed through torch.softmax / argmax."""
    from b200vit.inference import GraphedForward, predict
    from b200vit.models import vit
    from common import load_npz, sub
    from oracle import vit_oracle as O
    tiny = load_npz("vit_tiny_train.npz")
    sd = sub(tiny, "sd0/")
    model = vit.ViT(**TINY)
    model.load_state_dict(sd)
    X = torch.from_numpy(tiny["X"])
    ref_logits = O.vit_forward(X, sd, O.ViTConfig(**TINY))
    ref_probs = torch.softmax(ref_logits, dim=1)
    probs, labels = predict(model, X, device=_dev())
    assert not model.training and next(model.parameters()).is_cuda
    assert probs.shape == ref_probs.shape and labels.dtype == torch.int64
    assert rel_err(probs, ref_probs) < 2e-2                                   # bf16-mode tolerance on the logits
    assert torch.allclose(probs.sum(dim=1), torch.ones(len(X), device=_dev()), atol=1e-5)
    assert torch.equal(labels.cpu(), probs.argmax(dim=1).cpu())
    # labels agree with the oracle wherever its top two classes are not a near-tie
    top2 = ref_probs.topk(2, dim=1).values
    clear = (top2[:, 0] - top2[:, 1]) > 0.02
    assert torch.equal(labels.cpu()[clear], ref_probs.argmax(dim=1)[clear])
    fwd = GraphedForward(model)
    for _ in range(2):                                                        # capture, then replay
        p2, l2 = predict(model, X[:1], graphed=fwd)
        assert rel_err(p2, ref_probs[:1]) < 2e-2 and l2.shape == (1,)
    # the graph replay against the oracle directly (not only against the eager forward)
    assert rel_err(fwd(X.to(_dev())), ref_logits) < 2e-2
    # uint8 images go straight in
    xu8 = (X * 255).to(torch.uint8)
    p3, _ = predict(model, xu8)
    assert rel_err(p3, torch.softmax(O.vit_forward(xu8.float() / 255, sd, O.ViTConfig(**TINY)), dim=1)) < 2e-2
