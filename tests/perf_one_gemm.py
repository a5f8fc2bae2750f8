"""Launch the two representative GEMMs of the step a few times (for ncu captures; not a pytest file)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
from b200vit import kernels as K
dev = torch.device("cuda:0")
R, D, M = 50432, 768, 3072
ln = torch.randn(R, D, device=dev).bfloat16(); w1 = torch.randn(M, D, device=dev).bfloat16(); b1 = torch.randn(M, device=dev)
a = torch.empty(R, M, device=dev, dtype=torch.bfloat16); h = torch.empty_like(a)
da = torch.randn(R, M, device=dev).bfloat16(); dln = torch.empty(R, D, device=dev, dtype=torch.bfloat16)
for _ in range(3):
    K.gemm(K.OP_NT, K.EPI_BIAS_GELU, R, M, D, ln, D, w1, D, a, M, C2=h, ldc2=M, bias=b1, dropout_p=0.1, seed=1, offset=2)   # fc1 forward
    K.gemm(K.OP_NN, K.EPI_BF16, R, D, M, da, M, w1, D, dln, D)                                                            # fc1 dgrad
torch.cuda.synchronize()
print("ok")
