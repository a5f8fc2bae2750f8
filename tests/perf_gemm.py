"""Quick GEMM throughput probe (not a pytest file): python tests/perf_gemm.py"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
from b200vit import kernels as K

dev = torch.device("cuda:0")
R = 50432
shapes = [  # (name, op, epi, M, N, Kd, split)
    ("qkv_fwd", K.OP_NT, K.EPI_BF16, R, 2304, 768, 1),
    ("proj_fwd", K.OP_NT, K.EPI_BIAS_RES, R, 768, 768, 1),
    ("fc1_fwd", K.OP_NT, K.EPI_BIAS_GELU, R, 3072, 768, 1),
    ("fc2_fwd", K.OP_NT, K.EPI_BIAS_RES, R, 768, 3072, 1),
    ("fc2_dgrad", K.OP_NN, K.EPI_DGELU, R, 3072, 768, 1),
    ("fc1_dgrad", K.OP_NN, K.EPI_BF16, R, 768, 3072, 1),
    ("qkv_dgrad", K.OP_NN, K.EPI_BF16, R, 768, 2304, 1),
    ("fc1_wgrad", K.OP_TN, K.EPI_ACCUM, 3072, 768, R, 0),
    ("fc2_wgrad", K.OP_TN, K.EPI_ACCUM, 768, 3072, R, 0),
    ("qkv_wgrad", K.OP_TN, K.EPI_ACCUM, 2304, 768, R, 0),
    ("proj_wgrad", K.OP_TN, K.EPI_ACCUM, 768, 768, R, 0),
]
for name, op, epi, M, N, Kd, split in shapes:
    if op == K.OP_NT:
        A = torch.randn(M, Kd, device=dev).bfloat16(); B = torch.randn(N, Kd, device=dev).bfloat16(); lda, ldb = Kd, Kd
    elif op == K.OP_NN:
        A = torch.randn(M, Kd, device=dev).bfloat16(); B = torch.randn(Kd, N, device=dev).bfloat16(); lda, ldb = Kd, N
    else:
        A = torch.randn(Kd, M, device=dev).bfloat16(); B = torch.randn(Kd, N, device=dev).bfloat16(); lda, ldb = M, N
    f32 = epi in (K.EPI_BIAS_RES, K.EPI_ACCUM)
    C = torch.zeros(M, N, device=dev, dtype=torch.float32 if f32 else torch.bfloat16)
    C2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16) if epi == K.EPI_BIAS_GELU else None
    bias = torch.randn(N, device=dev) if epi in (K.EPI_BF16, K.EPI_BIAS_GELU, K.EPI_BIAS_RES) and op == K.OP_NT else None
    aux = None; ldaux = 0
    if epi == K.EPI_BIAS_RES: aux = torch.randn(M, N, device=dev); ldaux = N
    if epi == K.EPI_DGELU: aux = torch.randn(M, N, device=dev).bfloat16(); ldaux = N
    if split == 0: split = 0
    for p in (0.0, 0.1) if epi in (K.EPI_BIAS_GELU, K.EPI_BIAS_RES, K.EPI_DGELU) else (0.0,):
        def run():
            K.gemm(op, epi, M, N, Kd, A, lda, B, ldb, C, N, C2=C2, ldc2=N, bias=bias, aux=aux, ldaux=ldaux,
                   dropout_p=p, seed=1, offset=2, split_k=split)
        for _ in range(3): run()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        n = 10
        e0.record()
        for _ in range(n): run()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(f"{name:12s} p={p:.1f} split={split} {ms*1e3:8.1f} us  {2*M*N*Kd/ms/1e9:8.1f} TFLOP/s", flush=True)
# torch reference for scale
A = torch.randn(R, 768, device=dev).bfloat16(); B = torch.randn(3072, 768, device=dev).bfloat16()
for _ in range(3): A @ B.t()
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): A @ B.t()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"cublas fc1 plain {ms*1e3:8.1f} us {2*R*3072*768/ms/1e9:8.1f} TFLOP/s")
