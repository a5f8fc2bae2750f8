"""Correctness AT the benchmarked shapes (BASELINE.json configs[1]: ViT-B/16, batch 256 -> 50 432 token rows):
every GEMM launch shape of one training step (bench.gemm_shapes_per_step(256): 394 M-tiles, multi-wave persistent
2-CTA scheduling, automatic split-K) with its real epilogue against torch.matmul in fp32 (TF32 off) on the same bf16
operands, and one full forward + backward of the model at batch 256 against the oracle executed on the same GPU."""
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from common import rel_err  # noqa: E402
from oracle import vit_oracle as O  # noqa: E402

pytestmark = pytest.mark.gpu
BATCH = 256


def dev():
    return torch.device("cuda:0")


def _shapes():
    import bench
    return bench.gemm_shapes_per_step(BATCH)


@pytest.mark.parametrize("idx", range(14))
def test_gemm_at_bench_shape(idx):
    from b200vit import kernels as K
    torch.backends.cuda.matmul.allow_tf32 = False
    name, op, epi, M, N, Kd, _count = _shapes()[idx]
    g = torch.Generator(device="cuda").manual_seed(idx)
    rnd = lambda *s, scale=1.0: (torch.randn(*s, generator=g, device=dev()) * scale)
    if op == K.OP_NT:
        A, B = rnd(M, Kd).bfloat16(), rnd(N, Kd, scale=0.05).bfloat16()
        acc = A.float() @ B.float().t()
        lda, ldb = Kd, Kd
    elif op == K.OP_NN:
        A, B = rnd(M, Kd).bfloat16(), rnd(Kd, N, scale=0.05).bfloat16()
        acc = A.float() @ B.float()
        lda, ldb = Kd, N
    else:
        A, B = rnd(Kd, M, scale=0.05).bfloat16(), rnd(Kd, N, scale=0.05).bfloat16()
        acc = A.float().t() @ B.float()
        lda, ldb = M, N
    bias = rnd(N) if op == K.OP_NT else None
    if epi == K.EPI_PATCH:
        pos = rnd(197, N)
        C = torch.zeros(M + M // 196, N, device=dev())
        K.gemm(op, epi, M, N, Kd, A, lda, B, ldb, C, N, bias=bias, aux=pos, ldaux=N, row_period=196)
        ref = (acc + bias).view(M // 196, 196, N) + pos[1:]
        got = C.view(M // 196, 197, N)
        assert rel_err(got[:, 1:], ref) < 1e-4, name
        assert (got[:, 0] == 0).all(), "class-token rows must be left alone"
    elif epi == K.EPI_BF16:
        C = torch.empty(M, N, dtype=torch.bfloat16, device=dev())
        K.gemm(op, epi, M, N, Kd, A, lda, B, ldb, C, N, bias=bias)
        assert rel_err(C, acc + bias if bias is not None else acc) < 4e-3, name     # bf16 output rounding
    elif epi == K.EPI_BIAS_RES:
        res = rnd(M, N)
        C = torch.empty(M, N, device=dev())
        K.gemm(op, epi, M, N, Kd, A, lda, B, ldb, C, N, bias=bias, aux=res, ldaux=N)
        assert rel_err(C, acc + bias + res) < 1e-4, name
    elif epi == K.EPI_BIAS_GELU:
        gd = torch.empty(M, N, dtype=torch.bfloat16, device=dev())
        h = torch.empty(M, N, dtype=torch.bfloat16, device=dev())
        K.gemm(op, epi, M, N, Kd, A, lda, B, ldb, gd, N, C2=h, ldc2=N, bias=bias)
        a = acc + bias
        assert rel_err(h, torch.nn.functional.gelu(a)) < 5e-3, name
        cdf = 0.5 * (1 + torch.erf(a * 0.7071067811865476))
        assert rel_err(gd, cdf + a * torch.exp(-0.5 * a * a) * 0.3989422804014327) < 5e-3, name
    elif epi == K.EPI_DGELU:
        aux = rnd(M, N).bfloat16()
        C = torch.empty(M, N, dtype=torch.bfloat16, device=dev())
        colsum = torch.zeros(N, device=dev())
        K.gemm(op, epi, M, N, Kd, A, lda, B, ldb, C, N, bias=colsum, aux=aux, ldaux=N)
        assert rel_err(C, acc * aux.float()) < 5e-3, name
        assert rel_err(colsum, C.float().sum(0)) < 2e-3, name
    else:   # EPI_ACCUM, automatic split-K
        C = torch.ones(M, N, device=dev())
        K.gemm(op, epi, M, N, Kd, A, lda, B, ldb, C, N, split_k=0)
        assert rel_err(C, acc + 1.0) < 1e-4, name


def test_vit_b16_batch256_forward_backward_vs_oracle_on_gpu():
    """The bench workload itself (dropout off so that it is comparable): loss, logits and the gradient norm of every
    parameter tensor against the oracle run in fp32 on the same device."""
    from b200vit.models.vit import ViT
    from b200vit.nn import CrossEntropyLoss
    torch.backends.cuda.matmul.allow_tf32 = False
    kw = dict(num_classes=3, mlp_dropout=0.0, embedding_dropout=0.0)
    torch.manual_seed(42)
    m = ViT(**kw)
    sd = {k: v.clone().to(dev()).requires_grad_(True) for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(1234)
    X = torch.rand(BATCH, 3, 224, 224, generator=g).to(dev())
    y = torch.randint(0, 3, (BATCH,), generator=g).to(dev())
    # oracle in 4 slices of 64 images (the T x T scores of 256 images at once are 1.4 GB per layer); gradients add up
    ref_logits, ref_loss = [], 0.0
    for i in range(0, BATCH, 64):
        lg = O.vit_forward(X[i:i + 64], sd, O.ViTConfig(**kw))
        part = O.cross_entropy(lg, y[i:i + 64]) * (64 / BATCH)
        part.backward()
        ref_logits.append(lg.detach())
        ref_loss += part.item()
    ref_logits = torch.cat(ref_logits)
    m = m.to(dev()).train()
    logits = m(X)
    loss = CrossEntropyLoss()(logits, y)
    loss.backward()
    assert rel_err(logits, ref_logits) < 2e-2                               # north_star bf16 tolerance
    assert abs(loss.item() - ref_loss) < 1e-2 * ref_loss
    tot_ours = torch.sqrt(sum(p.grad.double().pow(2).sum() for p in m.parameters())).item()
    tot_ref = torch.sqrt(sum(v.grad.double().pow(2).sum() for v in sd.values())).item()
    assert abs(tot_ours - tot_ref) < 2e-2 * tot_ref, (tot_ours, tot_ref)
    worst = {}
    for k, p in m.named_parameters():
        if k.endswith("in_proj_bias"):
            continue
        worst[k] = rel_err(p.grad, sd[k].grad)
    bad = {k: v for k, v in worst.items() if v > 6e-2}      # gradient tolerance of the bf16 path
    assert not bad, bad
