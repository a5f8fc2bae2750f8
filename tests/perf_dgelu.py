"""fc2 dgrad (dGELU epilogue) with and without the fused fc1 bias gradient, against the plain bf16 epilogue on the same
shape (not a pytest file). Measured: 214 us / 244 us / 177 us."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200")); sys.path.insert(0, str(ROOT))
import torch
from b200vit import kernels as K
dev = torch.device("cuda:0")
M, N, Kd = 50432, 3072, 768
A = torch.randn(M, Kd, device=dev).bfloat16(); B = torch.randn(Kd, N, device=dev).bfloat16()
C = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
aux = torch.randn(M, N, device=dev).bfloat16()
dbias = torch.zeros(N, device=dev)
def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
print("dGELU, no bias gradient :", round(timed(lambda: K.gemm(K.OP_NN, K.EPI_DGELU, M, N, Kd, A, Kd, B, N, C, N, aux=aux, ldaux=N)), 1), "us")
print("dGELU + bias gradient   :", round(timed(lambda: K.gemm(K.OP_NN, K.EPI_DGELU, M, N, Kd, A, Kd, B, N, C, N, aux=aux, ldaux=N, bias=dbias)), 1), "us")
print("plain bf16 epilogue     :", round(timed(lambda: K.gemm(K.OP_NN, K.EPI_BF16, M, N, Kd, A, Kd, B, N, C, N)), 1), "us")
