"""LayerNorm folded into the GEMMs around it (forward passes without a backward; csrc/gemm_sm100.cu epilogues 7-10,
b200vit.functional "LayerNorm folded ..."): kernel-level checks against torch, model-level parity against the oracle with
the fold on and off (reference models/vit.py:93-97, 129-130: LayerNorm feeding in_proj / fc1)."""
import pytest
import torch

from common import TINY, load_npz, rel_err, sub
from oracle import vit_oracle as O

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


@pytest.mark.parametrize("M,N,Kd", [(384, 768, 512), (6304, 768, 768), (200, 328, 264)])
def test_producer_epilogue_writes_bf16_copy_and_row_statistics(M, N, Kd):
    from b200vit import _C, kernels as K
    g = torch.Generator().manual_seed(M)
    A = (torch.randn(M, Kd, generator=g)).bfloat16().to(dev())
    W = (torch.randn(N, Kd, generator=g) * 0.05).bfloat16().to(dev())
    bias = torch.randn(N, generator=g).to(dev())
    res = (torch.randn(M, N, generator=g) + 0.5).to(dev())
    C = torch.empty(M, N, device=dev())
    x16 = torch.empty(M, N, dtype=torch.bfloat16, device=dev())
    stats = torch.zeros(M, 2, device=dev())
    K.gemm_ln(_C.EPI_RES_STATS, M, N, Kd, A, Kd, W, Kd, C, N, bias=bias, aux=res, ldaux=N, x16=x16, ldx16=N, rowstats=stats)
    ref = A.float() @ W.float().t() + bias + res
    assert rel_err(C, ref) < 1e-4
    assert torch.equal(x16, C.bfloat16())
    assert rel_err(stats[:, 0], C.double().sum(1)) < 1e-5
    assert rel_err(stats[:, 1], (C.double() ** 2).sum(1)) < 1e-5


@pytest.mark.parametrize("M,N,D", [(384, 768, 256), (6304, 2304, 768), (200, 328, 264)])
def test_consumer_epilogues_equal_layernorm_then_linear(M, N, D):
    from b200vit import _C, kernels as K
    g = torch.Generator().manual_seed(N)
    x = (torch.randn(M, D, generator=g) * 0.7 + 0.5).to(dev())            # mean 0.5, as the reference's residual stream
    gamma = (1 + 0.2 * torch.randn(D, generator=g)).to(dev())
    beta = (0.1 * torch.randn(D, generator=g)).to(dev())
    W = (torch.randn(N, D, generator=g) * 0.05).to(dev())
    b = torch.randn(N, generator=g).to(dev())
    ref = torch.nn.functional.layer_norm(x, (D,), gamma, beta, 1e-5) @ W.t() + b
    w16 = torch.empty(N, D, dtype=torch.bfloat16, device=dev())
    s = torch.empty(N, device=dev())
    c = torch.empty(N, device=dev())
    K.fold_ln_weight(W, gamma, beta, b, w16, s, c)
    assert torch.equal(w16, (W * gamma).bfloat16())
    assert rel_err(s, w16.float().sum(1)) < 1e-5 and rel_err(c, (W * beta).sum(1) + b) < 1e-5
    stats = torch.stack((x.sum(1), (x * x).sum(1)), dim=1).contiguous()
    out = torch.empty(M, N, dtype=torch.bfloat16, device=dev())
    K.gemm_ln(_C.EPI_LN_BF16, M, N, D, x.bfloat16(), D, w16, D, out, N, bias=c, ln_stats=stats, ln_s=s, ln_dim=D, ln_eps=1e-5)
    # the unfolded bf16 path for scale: LayerNorm output rounded to bf16, then the GEMM
    plain = torch.nn.functional.layer_norm(x, (D,), gamma, beta, 1e-5).bfloat16().float() @ W.bfloat16().float().t() + b
    e_fold, e_plain = rel_err(out, ref), rel_err(plain.bfloat16(), ref)
    assert e_fold < 1e-2 and e_fold < 3 * e_plain + 1e-3, (e_fold, e_plain)
    h = torch.empty(M, N, dtype=torch.bfloat16, device=dev())
    K.gemm_ln(_C.EPI_LN_GELU, M, N, D, x.bfloat16(), D, w16, D, None, N, C2=h, ldc2=N, bias=c, ln_stats=stats, ln_s=s, ln_dim=D,
              ln_eps=1e-5)
    assert rel_err(h, torch.nn.functional.gelu(ref)) < 1.5e-2


def test_model_forward_with_and_without_the_fold():
    """Eval forward under inference_mode: fold on (default) and off, both within the bf16 tolerance of the reference."""
    import b200vit.functional as F_
    from b200vit import _C
    from b200vit.models.vit import ViT
    F_.LN_FOLD_MIN_ROWS, keep_rows = 0, F_.LN_FOLD_MIN_ROWS      # the fold is for large batches by default: force it here
    request_restore = keep_rows
    tiny = load_npz("vit_tiny_train.npz")
    m = ViT(**TINY)
    m.load_state_dict(sub(tiny, "sd0/"))
    m = m.to(dev()).eval()
    X = torch.from_numpy(tiny["X"]).to(dev())
    golden = torch.from_numpy(tiny["logits_eval"])
    outs = {}
    for flag in (True, False):
        prev = F_.set_ln_fold(flag)
        try:
            c0 = _C.call_count
            with torch.inference_mode():
                outs[flag] = m(X).clone()
            calls = _C.call_count - c0
        finally:
            F_.set_ln_fold(prev)
        assert rel_err(outs[flag], golden) < 2e-2, (flag, rel_err(outs[flag], golden))
        outs[(flag, "calls")] = calls
    assert rel_err(outs[True], outs[False]) < 1e-2
    assert not torch.equal(outs[True], outs[False])                 # the folded path really ran
    # with grad enabled and trainable parameters the LayerNorm kernels are back (the backward needs their output)
    m.train()
    loss = m(X).square().mean()
    loss.backward()
    assert all(p.grad is not None for p in m.parameters())
    F_.LN_FOLD_MIN_ROWS = request_restore


def test_vit_b16_inference_logits_with_fold_vs_oracle():
    import b200vit.functional as F_
    from b200vit.models.vit import ViT
    torch.manual_seed(42)
    m = ViT(num_classes=3, mlp_dropout=0.0, embedding_dropout=0.0)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(0)
    X = torch.rand(4, 3, 224, 224, generator=g)
    ref = O.vit_forward(X, sd, O.ViTConfig(num_classes=3, mlp_dropout=0.0, embedding_dropout=0.0))
    m = m.to(dev()).eval()
    assert F_._LN_FOLD
    keep_rows, F_.LN_FOLD_MIN_ROWS = F_.LN_FOLD_MIN_ROWS, 0
    try:
        with torch.inference_mode():
            out = m(X.to(dev()))
    finally:
        F_.LN_FOLD_MIN_ROWS = keep_rows
    err = rel_err(out, ref)
    print(f"ViT-B/16 inference logits rel err with the LayerNorm fold: {err:.2e}")
    assert err < 2e-2, err
