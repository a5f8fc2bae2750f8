"""One launch of every hot kernel of the ViT-B/16 B=256 training step, for `ncu --set full` (not a pytest file).

    python tools/profile_kernels.py > gpurun_out/plain.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:'gemm_tcgen05|attention|layernorm|colsum|dropout_cast|adam' \
        -o gpurun_out/r01_kernels python tools/profile_kernels.py
    ncu -i gpurun_out/r01_kernels.ncu-rep --page raw --csv > /tmp/raw.csv; python tools/summarize_ncu.py /tmp/raw.csv

Launch order (the summariser relies on it): the 14 GEMM shapes of bench.gemm_shapes_per_step in order, then
attention forward, attention backward, LayerNorm forward, LayerNorm backward.
"""
import math
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
sys.path.insert(0, str(ROOT))
import torch

import bench
from b200vit import kernels as K

B = 256
dev = torch.device("cuda:0")
torch.manual_seed(0)
for name, op, epi, M, N, Kd, count in bench.gemm_shapes_per_step(B):
    run = bench.make_gemm_runner(name, op, epi, M, N, Kd, dev)
    run()
    torch.cuda.synchronize()
    del run

T, H, hd, D = 197, 12, 64, 768
qkv = torch.randn(B, T, 3, H, hd, device=dev).bfloat16()
out = torch.empty(B, T, H * hd, device=dev, dtype=torch.bfloat16)
lse = torch.empty(B, H, T, device=dev)
dout = torch.randn(B, T, H * hd, device=dev).bfloat16()
delta = torch.empty(B, H, T, device=dev)
dqkv = torch.empty_like(qkv)
sc = 1 / math.sqrt(hd)
K.attention_fwd(qkv, out, lse, B, T, H, hd, sc)
K.attention_bwd(qkv, out, dout, lse, delta, dqkv, B, T, H, hd, sc)
torch.cuda.synchronize()

# LayerNorm forward / backward through the autograd function the model uses
from b200vit import functional as F_
x = torch.randn(B * T, D, device=dev, requires_grad=True)
g = torch.ones(D, device=dev, requires_grad=True)
b = torch.zeros(D, device=dev, requires_grad=True)
y = F_.LayerNormFn.apply(x, g, b)
y.backward(torch.randn_like(y))
torch.cuda.synchronize()
print("ok")
