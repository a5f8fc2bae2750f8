"""Per-kernel parity: each C-ABI entry point on the GPU vs a plain torch fp32 reference of the same op
(floating-point kernels; tolerances are written next to each check)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from b200vit import kernels as K  # noqa: E402


def dev():
    return torch.device("cuda:0")


def rel_err(a, b):
    a = a.float()
    b = b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _rand_bf16(shape, gen, scale=1.0):
    return (torch.randn(shape, generator=gen, device="cpu") * scale).to(torch.bfloat16).to(dev())


@pytest.mark.parametrize("M,N,Kd", [(128, 256, 64), (256, 512, 768), (200, 328, 200), (1024, 768, 3072), (6304, 2304, 768)])
def test_gemm_nt_bias(M, N, Kd):
    g = torch.Generator().manual_seed(1)
    A = _rand_bf16((M, Kd), g)
    B = _rand_bf16((N, Kd), g, 0.05)
    bias = torch.randn(N, generator=g).to(dev())
    C = torch.empty((M, N), dtype=torch.bfloat16, device=dev())
    K.gemm(K.OP_NT, K.EPI_BF16, M, N, Kd, A, Kd, B, Kd, C, N, bias=bias)
    ref = A.float() @ B.float().t() + bias
    # bf16 output rounding: 2^-9 relative per element
    assert rel_err(C, ref) < 4e-3
    assert (C.float() - ref).abs().max().item() < 0.05 * ref.abs().max().item() + 1e-2


def test_gemm_dynamic_and_static_work_distribution_agree():
    """b200vit_set_gemm_schedule: work items claimed from the device counter (default) or split statically give
    bit-identical results (an item is computed the same way whichever cluster takes it), launch after launch -- the
    counter is put back to zero by every launch."""
    from b200vit import _C
    g = torch.Generator().manual_seed(11)
    M, N, Kd = 12608, 3072, 768                        # 50 x 12 = 600 work items for 74 clusters
    A = _rand_bf16((M, Kd), g)
    B = _rand_bf16((N, Kd), g, 0.05)
    bias = torch.randn(N, generator=g).to(dev())
    outs = {}
    try:
        for mode in (True, False, True):
            _C.set_gemm_schedule(mode)
            for rep in range(40):
                C = torch.zeros((M, N), dtype=torch.bfloat16, device=dev())
                K.gemm(K.OP_NT, K.EPI_BF16, M, N, Kd, A, Kd, B, Kd, C, N, bias=bias)
                if mode in outs:
                    assert torch.equal(C, outs[mode]), f"dynamic={mode}: launch {rep} differs from the first"
                else:
                    outs[mode] = C
    finally:
        _C.set_gemm_schedule(True)
    assert torch.equal(outs[True], outs[False])
    assert rel_err(outs[True], A.float() @ B.float().t() + bias) < 4e-3


@pytest.mark.parametrize("M,N,Kd", [(256, 256, 128), (200, 328, 200), (6304, 768, 2304)])
def test_gemm_nn(M, N, Kd):
    g = torch.Generator().manual_seed(2)
    A = _rand_bf16((M, Kd), g)
    B = _rand_bf16((Kd, N), g, 0.05)
    C = torch.empty((M, N), dtype=torch.bfloat16, device=dev())
    K.gemm(K.OP_NN, K.EPI_BF16, M, N, Kd, A, Kd, B, N, C, N)
    ref = A.float() @ B.float()
    assert rel_err(C, ref) < 4e-3


@pytest.mark.parametrize("M,N,Kd,split", [(256, 256, 128, 1), (200, 328, 264, 2), (768, 3072, 6304, 8), (2304, 768, 6304, 5)])
def test_gemm_tn_accum(M, N, Kd, split):
    g = torch.Generator().manual_seed(3)
    A = _rand_bf16((Kd, M), g)
    B = _rand_bf16((Kd, N), g, 0.05)
    C = torch.ones((M, N), dtype=torch.float32, device=dev())
    K.gemm(K.OP_TN, K.EPI_ACCUM, M, N, Kd, A, M, B, N, C, N, split_k=split)
    ref = A.float().t() @ B.float() + 1.0
    assert rel_err(C, ref) < 1e-4


def test_gemm_gelu_and_dgelu():
    g = torch.Generator().manual_seed(4)
    M, N, Kd = 512, 768, 256
    A = _rand_bf16((M, Kd), g)
    B = _rand_bf16((N, Kd), g, 0.08)
    bias = torch.randn(N, generator=g).to(dev()) * 0.1
    gd = torch.empty((M, N), dtype=torch.bfloat16, device=dev())    # gelu'(a): what the backward epilogue multiplies by
    h = torch.empty((M, N), dtype=torch.bfloat16, device=dev())
    K.gemm(K.OP_NT, K.EPI_BIAS_GELU, M, N, Kd, A, Kd, B, Kd, gd, N, C2=h, ldc2=N, bias=bias)
    a_ref = (A.float() @ B.float().t() + bias).requires_grad_(True)
    h_ref = torch.nn.functional.gelu(a_ref)
    h_ref.sum().backward()
    assert rel_err(h, h_ref) < 5e-3            # erf via A&S 7.1.25 (2.5e-5) + bf16 output rounding
    assert rel_err(gd, a_ref.grad) < 5e-3
    # backward epilogue: out = (G @ W) * gelu'(a)
    G = _rand_bf16((M, Kd), g)
    W = _rand_bf16((Kd, N), g, 0.08)
    out = torch.empty((M, N), dtype=torch.bfloat16, device=dev())
    K.gemm(K.OP_NN, K.EPI_DGELU, M, N, Kd, G, Kd, W, N, out, N, aux=gd, ldaux=N)
    assert rel_err(out, (G.float() @ W.float()) * a_ref.grad) < 8e-3
    # ... and, optionally, the column sums of the same product (the fc1 bias gradient) accumulated into `bias`
    out2 = torch.empty_like(out)
    colsum = torch.full((N,), 1.0, dtype=torch.float32, device=dev())       # accumulates on top of what is there
    K.gemm(K.OP_NN, K.EPI_DGELU, M, N, Kd, G, Kd, W, N, out2, N, bias=colsum, aux=gd, ldaux=N)
    assert torch.equal(out2, out)
    assert rel_err(colsum - 1.0, ((G.float() @ W.float()) * gd.float()).sum(0)) < 2e-3
    # ragged shape (edge tiles)
    Mr, Nr, Kr = 328, 200, 136
    Gr, Wr = _rand_bf16((Mr, Kr), g), _rand_bf16((Kr, Nr), g, 0.08)
    gdr = _rand_bf16((Mr, Nr), g)
    outr = torch.empty((Mr, Nr), dtype=torch.bfloat16, device=dev())
    csr = torch.zeros(Nr, dtype=torch.float32, device=dev())
    K.gemm(K.OP_NN, K.EPI_DGELU, Mr, Nr, Kr, Gr, Kr, Wr, Nr, outr, Nr, bias=csr, aux=gdr, ldaux=Nr)
    assert rel_err(outr, (Gr.float() @ Wr.float()) * gdr.float()) < 8e-3
    assert rel_err(csr, ((Gr.float() @ Wr.float()) * gdr.float()).sum(0)) < 3e-3
    # with dropout: hidden and saved derivative carry the SAME mask, scaled by 1/(1-p); mask == dropout_cast's
    p = 0.25
    gd2, h2 = torch.empty_like(gd), torch.empty_like(h)
    K.gemm(K.OP_NT, K.EPI_BIAS_GELU, M, N, Kd, A, Kd, B, Kd, gd2, N, C2=h2, ldc2=N, bias=bias, dropout_p=p, seed=9, offset=3)
    ones = torch.ones((M, N), dtype=torch.float32, device=dev())
    mask = torch.empty((M, N), dtype=torch.bfloat16, device=dev())
    K.dropout_cast(ones, mask, None, M, N, p, 9, 3)
    keep = mask != 0
    assert abs(keep.float().mean().item() - (1 - p)) < 0.01
    assert torch.all(h2[~keep] == 0) and torch.all(gd2[~keep] == 0)
    assert rel_err(h2[keep], (h_ref.detach() / (1 - p))[keep]) < 5e-3
    assert rel_err(gd2[keep], (a_ref.grad / (1 - p))[keep]) < 5e-3


def test_gemm_bias_res_and_dropout_consistency():
    g = torch.Generator().manual_seed(5)
    M, N, Kd = 384, 768, 512
    A = _rand_bf16((M, Kd), g)
    B = _rand_bf16((N, Kd), g, 0.05)
    bias = torch.randn(N, generator=g).to(dev())
    res = torch.randn((M, N), generator=g).to(dev())
    C = torch.empty((M, N), dtype=torch.float32, device=dev())
    K.gemm(K.OP_NT, K.EPI_BIAS_RES, M, N, Kd, A, Kd, B, Kd, C, N, bias=bias, aux=res, ldaux=N)
    ref = A.float() @ B.float().t() + bias + res
    assert rel_err(C, ref) < 1e-4
    # dropout: same (seed, offset) -> same mask in the GEMM epilogue and in dropout_cast
    p = 0.25
    C2 = torch.empty_like(C)
    K.gemm(K.OP_NT, K.EPI_BIAS_RES, M, N, Kd, A, Kd, B, Kd, C2, N, bias=bias, dropout_p=p, seed=123, offset=7)
    y = A.float() @ B.float().t() + bias
    keep = (C2 != 0)
    frac = keep.float().mean().item()
    assert abs(frac - (1 - p)) < 0.01
    assert rel_err(C2[keep], (y / (1 - p))[keep]) < 1e-4
    ones = torch.ones((M, N), dtype=torch.float32, device=dev())
    mask_bf16 = torch.empty((M, N), dtype=torch.bfloat16, device=dev())
    colsum = torch.empty(N, dtype=torch.float32, device=dev())
    K.dropout_cast(ones, mask_bf16, colsum, M, N, p, 123, 7)
    assert torch.equal(mask_bf16 != 0, keep)
    assert rel_err(colsum, mask_bf16.float().sum(0)) < 1e-5


def test_patch_embedding_pieces():
    g = torch.Generator().manual_seed(6)
    Bn, Cc, H, P, D = 3, 3, 64, 16, 128
    img = torch.rand((Bn, Cc, H, H), generator=g).to(dev())
    w = (torch.randn((D, Cc, P, P), generator=g) * 0.05).to(dev())
    bias = torch.randn(D, generator=g).to(dev())
    Np = (H // P) ** 2
    T = Np + 1
    cls = torch.rand((1, 1, D), generator=g).to(dev())
    pos = torch.rand((1, T, D), generator=g).to(dev())
    patches = K.im2col_patches(img, P)
    ref_patches = torch.nn.functional.unfold(img, P, stride=P).transpose(1, 2).reshape(Bn * Np, -1)
    assert torch.equal(patches.float(), ref_patches.to(torch.bfloat16).float())
    x = torch.full((Bn * T, D), float("nan"), dtype=torch.float32, device=dev())
    wb = K.cast_bf16(w.reshape(D, -1).contiguous())
    Kd = Cc * P * P
    K.gemm(K.OP_NT, K.EPI_PATCH, Bn * Np, D, Kd, patches, Kd, wb, Kd, x, D, bias=bias, aux=pos.reshape(T, D), ldaux=D,
           row_period=Np)
    K.class_token_rows(cls, pos, x, Bn, T, D, 0.0, 0, 0)
    conv = torch.nn.functional.conv2d(img, w, bias, stride=P).flatten(2).permute(0, 2, 1)
    ref = torch.cat((cls.expand(Bn, -1, -1), conv), dim=1) + pos
    assert rel_err(x.reshape(Bn, T, D), ref) < 3e-3


@pytest.mark.parametrize("rows,D", [(37, 64), (1000, 768), (300, 1024)])
def test_layernorm_fwd_bwd(rows, D):
    g = torch.Generator().manual_seed(7)
    x = (torch.randn((rows, D), generator=g) * 2 + 0.5).to(dev())
    gamma = (torch.randn(D, generator=g) * 0.2 + 1).to(dev())
    beta = torch.randn(D, generator=g).to(dev())
    y16 = torch.empty((rows, D), dtype=torch.bfloat16, device=dev())
    y32 = torch.empty((rows, D), dtype=torch.float32, device=dev())
    mean = torch.empty(rows, device=dev())
    rstd = torch.empty(rows, device=dev())
    K.layernorm_fwd(x, D, gamma, beta, y16, y32, mean, rstd, rows, D, 1e-5)
    xr = x.clone().requires_grad_(True)
    gr = gamma.clone().requires_grad_(True)
    br = beta.clone().requires_grad_(True)
    ref = torch.nn.functional.layer_norm(xr, (D,), gr, br, 1e-5)
    assert rel_err(y32, ref) < 1e-5          # fp32 kernel vs fp32 torch
    assert rel_err(y16, ref) < 4e-3          # bf16 rounding
    dy = torch.randn((rows, D), generator=g).to(dev())
    gin = torch.randn((rows, D), generator=g).to(dev())
    ref.backward(dy)
    gout = torch.empty_like(x)
    g16 = torch.empty((rows, D), dtype=torch.bfloat16, device=dev())
    dgamma = torch.empty(D, device=dev())
    dbeta = torch.empty(D, device=dev())
    dbn = torch.empty(D, device=dev())
    K.layernorm_bwd(None, dy, x, D, mean, rstd, gamma, gin, gout, D, g16, dgamma, dbeta, dbn, rows, D)
    assert rel_err(gout, xr.grad + gin) < 1e-4
    assert rel_err(dgamma, gr.grad) < 1e-4
    assert rel_err(dbeta, br.grad) < 1e-4
    assert rel_err(g16, xr.grad + gin) < 4e-3
    assert rel_err(dbn, g16.float().sum(0)) < 1e-4
    # bf16 dy input path
    dy16 = dy.to(torch.bfloat16)
    K.layernorm_bwd(dy16, None, x, D, mean, rstd, gamma, None, gout, D, None, dgamma, dbeta, None, rows, D)
    xr.grad = None
    torch.nn.functional.layer_norm(xr, (D,), gamma, beta, 1e-5).backward(dy16.float())
    assert rel_err(gout, xr.grad) < 1e-4


@pytest.mark.parametrize("Bn,T,H", [(2, 197, 12), (1, 64, 2), (3, 50, 4), (2, 128, 3), (2, 129, 2), (1, 256, 2), (2, 240, 1), (7, 16, 1),
                                    (1, 577, 16), (2, 257, 3), (1, 1000, 2), (2, 384, 5), (1, 640, 1)])
def test_attention_fwd_bwd(Bn, T, H):
    g = torch.Generator().manual_seed(8)
    hd = 64
    qkv = _rand_bf16((Bn, T, 3, H, hd), g, 1.0)
    out = torch.empty((Bn, T, H * hd), dtype=torch.bfloat16, device=dev())
    lse = torch.empty((Bn, H, T), dtype=torch.float32, device=dev())
    scale = 1.0 / math.sqrt(hd)
    K.attention_fwd(qkv, out, lse, Bn, T, H, hd, scale)
    q, k, v = [qkv[:, :, i].float().permute(0, 2, 1, 3).detach().requires_grad_(True) for i in range(3)]
    s = (q @ k.transpose(-1, -2)) * scale
    ref = torch.softmax(s, -1) @ v
    ref_out = ref.permute(0, 2, 1, 3).reshape(Bn, T, H * hd)
    assert rel_err(out, ref_out) < 1e-2      # P is rounded to bf16 before the PV product
    assert rel_err(lse, torch.logsumexp(s, -1)) < 1e-4
    dout = _rand_bf16((Bn, T, H * hd), g, 1.0)
    delta = torch.empty((Bn, H, T), dtype=torch.float32, device=dev())
    dqkv = torch.empty_like(qkv)
    dbias = torch.full((3 * H * hd,), 7.0, dtype=torch.float32, device=dev())      # must be overwritten
    K.attention_bwd(qkv, out, dout, lse, delta, dqkv, Bn, T, H, hd, scale, dbias=dbias)
    assert rel_err(dbias, dqkv.float().reshape(Bn * T, -1).sum(0)) < 3e-3           # in_proj_bias gradient
    dqkv_plain = torch.empty_like(dqkv)
    K.attention_bwd(qkv, out, dout, lse, delta, dqkv_plain, Bn, T, H, hd, scale)
    # Same arithmetic with and without the fused bias gradient. Not bit-equal in general: the two MMA-issuing warps of
    # the kernel add their chunks' contributions to the dV / dK accumulators in a timing-dependent order (fp32
    # reassociation), which can move a value across a bf16 rounding boundary: at most one bf16 ulp, on a few elements.
    diff = (dqkv_plain.float() - dqkv.float()).abs()
    assert (diff <= dqkv.float().abs() * 2.0 ** -7 + 1e-30).all()
    assert (diff > 0).float().mean().item() < 1e-3
    ref.backward(dout.float().reshape(Bn, T, H, hd).permute(0, 2, 1, 3))
    for i, t in enumerate((q, k, v)):
        got = dqkv[:, :, i].float().permute(0, 2, 1, 3)
        assert rel_err(got, t.grad) < 2e-2, f"d{'qkv'[i]}"


@pytest.mark.parametrize("Bn,T,H,p", [(2, 197, 3, 0.1), (1, 70, 2, 0.5), (1, 300, 2, 0.2)])
def test_attention_probability_dropout(Bn, T, H, p):
    """nn.MultiheadAttention(dropout=p) semantics: softmax(S) is dropped (kept entries / (1-p)) before the PV product.
    The mask is regenerated here from the same (seed, offset) through dropout_cast on a [B*H*T, T] tensor of ones."""
    g = torch.Generator().manual_seed(12)
    hd = 64
    seed, offset = 77, 5
    qkv = _rand_bf16((Bn, T, 3, H, hd), g, 1.0)
    out = torch.empty((Bn, T, H * hd), dtype=torch.bfloat16, device=dev())
    lse = torch.empty((Bn, H, T), dtype=torch.float32, device=dev())
    scale = 1.0 / math.sqrt(hd)
    K.attention_fwd(qkv, out, lse, Bn, T, H, hd, scale, p, seed, offset)
    ones = torch.ones((Bn * H * T, T), dtype=torch.float32, device=dev())
    mask = torch.empty((Bn * H * T, T), dtype=torch.bfloat16, device=dev())
    K.dropout_cast(ones, mask, None, Bn * H * T, T, p, seed, offset)
    mask = mask.float().reshape(Bn, H, T, T)                      # 0 or 1/(1-p)
    assert abs((mask != 0).float().mean().item() - (1 - p)) < 0.01
    q, k, v = [qkv[:, :, i].float().permute(0, 2, 1, 3).detach().requires_grad_(True) for i in range(3)]
    s = (q @ k.transpose(-1, -2)) * scale
    ref = (torch.softmax(s, -1) * mask) @ v
    assert rel_err(out, ref.permute(0, 2, 1, 3).reshape(Bn, T, H * hd)) < 1e-2
    assert rel_err(lse, torch.logsumexp(s, -1)) < 1e-4            # the log-sum-exp is of the un-dropped scores
    dout = _rand_bf16((Bn, T, H * hd), g, 1.0)
    delta = torch.empty((Bn, H, T), dtype=torch.float32, device=dev())
    dqkv = torch.empty_like(qkv)
    dbias = torch.empty(3 * H * hd, dtype=torch.float32, device=dev())
    K.attention_bwd(qkv, out, dout, lse, delta, dqkv, Bn, T, H, hd, scale, dbias=dbias, dropout_p=p, seed=seed, offset=offset)
    ref.backward(dout.float().reshape(Bn, T, H, hd).permute(0, 2, 1, 3))
    for i, t in enumerate((q, k, v)):
        assert rel_err(dqkv[:, :, i].float().permute(0, 2, 1, 3), t.grad) < 2e-2, f"d{'qkv'[i]}"
    assert rel_err(dbias, dqkv.float().reshape(Bn * T, -1).sum(0)) < 3e-3
    # another (seed, offset) gives another mask
    out2 = torch.empty_like(out)
    K.attention_fwd(qkv, out2, lse, Bn, T, H, hd, scale, p, seed, offset + 1)
    assert rel_err(out2, out) > 0.05


@pytest.mark.parametrize("Bn,C", [(32, 3), (256, 1000), (7, 10)])
def test_head_and_ce(Bn, C):
    g = torch.Generator().manual_seed(9)
    D = 768
    xn = torch.randn((Bn, D), generator=g).to(dev())
    w = (torch.randn((C, D), generator=g) * 0.05).to(dev())
    b = torch.randn(C, generator=g).to(dev())
    logits = torch.empty((Bn, C), device=dev())
    K.head_fwd(xn, w, b, logits, Bn, D, C)
    ref = xn @ w.t() + b
    assert rel_err(logits, ref) < 1e-5
    y = torch.randint(0, C, (Bn,), generator=g).to(dev())
    loss = torch.empty(1, device=dev())
    dl = torch.empty((Bn, C), device=dev())
    correct = torch.empty(1, dtype=torch.int32, device=dev())
    K.softmax_ce(logits, y, loss, dl, correct, Bn, C)
    lr = ref.clone().requires_grad_(True)
    ref_loss = torch.nn.functional.cross_entropy(lr, y)
    ref_loss.backward()
    assert abs(loss.item() - ref_loss.item()) < 1e-5 * max(1.0, abs(ref_loss.item()))
    assert rel_err(dl, lr.grad) < 1e-5
    assert correct.item() == (ref.argmax(1) == y).sum().item()
    dxn = torch.empty_like(xn)
    dw = torch.empty_like(w)
    db = torch.empty_like(b)
    K.head_bwd(xn, w, dl, dxn, dw, db, Bn, D, C)
    assert rel_err(dxn, dl @ w) < 1e-5
    assert rel_err(dw, dl.t() @ xn) < 1e-5
    assert rel_err(db, dl.sum(0)) < 1e-5


def test_colsum_and_embed_grads():
    g = torch.Generator().manual_seed(10)
    rows, cols = 3000, 776
    x = _rand_bf16((rows, cols), g)
    out = torch.empty(cols, device=dev())
    K.colsum_bf16(x, cols, out, rows, cols)
    assert rel_err(out, x.float().sum(0)) < 1e-5
    Bn, T, D = 5, 17, 64
    gx = torch.randn((Bn, T, D), generator=g).to(dev())
    g_cls = torch.empty(D, device=dev())
    g_pos = torch.empty((T, D), device=dev())
    g_b = torch.empty(D, device=dev())
    K.embed_param_grads(gx, g_cls, g_pos, g_b, Bn, T, D, 0.0, 0, 0)
    assert rel_err(g_pos, gx.sum(0)) < 1e-6
    assert rel_err(g_cls, gx[:, 0].sum(0)) < 1e-6
    assert rel_err(g_b, gx[:, 1:].sum((0, 1))) < 1e-5
    gy = torch.empty((Bn * (T - 1), D), dtype=torch.bfloat16, device=dev())
    K.patch_grad_rows(gx, gy, Bn, T, D, 0.0, 0, 0)
    assert torch.equal(gy.float(), gx[:, 1:].reshape(-1, D).to(torch.bfloat16).float())
    # with the embedding dropout's mask: same mask as dropout_cast draws over the whole [B*T, D] tensor
    p = 0.3
    K.patch_grad_rows(gx, gy, Bn, T, D, p, 21, 4)
    full = torch.empty((Bn * T, D), dtype=torch.bfloat16, device=dev())
    K.dropout_cast(gx.reshape(Bn * T, D).contiguous(), full, None, Bn * T, D, p, 21, 4)
    assert torch.equal(gy, full.reshape(Bn, T, D)[:, 1:].reshape(-1, D))
