"""LayerNorm forward / backward alone at the training shape (50 432 x 768), the backward in its variants (not a pytest
file): python tests/perf_ln.py"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
from b200vit import kernels as K

dev = torch.device("cuda:0")
R, D = 256 * 197, 768
x = torch.randn(R, D, device=dev)
gamma = torch.rand(D, device=dev) + 0.5
beta = torch.randn(D, device=dev)
y16 = torch.empty(R, D, device=dev, dtype=torch.bfloat16)
mean = torch.empty(R, device=dev); rstd = torch.empty(R, device=dev)
dy16 = torch.randn(R, D, device=dev).to(torch.bfloat16)
gin = torch.randn(R, D, device=dev)
gout = torch.empty(R, D, device=dev)
g16 = torch.empty(R, D, device=dev, dtype=torch.bfloat16)
dg = torch.zeros(D, device=dev); db = torch.zeros(D, device=dev); dn = torch.zeros(D, device=dev)
big = torch.empty(64 * 1024 * 1024, device=dev)          # > L2, written between timed launches


def timed(fn, n=10):
    fn(); torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        big.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n * 1e3


us = timed(lambda: K.layernorm_fwd(x, D, gamma, beta, y16, None, mean, rstd, R, D, 1e-5))
print(f"fwd (f32 in, bf16 out): {us:.1f} us  {R * D * 6 / us / 1e6:.2f} TB/s")
for name, nxt, p, cols in (("plain (no next operand, no column sums)", False, 0.0, False),
                           ("column sums", False, 0.0, True),
                           ("next operand + bias gradient, no dropout", True, 0.0, True),
                           ("next operand + bias gradient, dropout 0.1 (the step's variant)", True, 0.1, True)):
    us = timed(lambda: K.layernorm_bwd(dy16, None, x, D, mean, rstd, gamma, gin, gout, D, g16 if nxt else None,
                                       dg if cols else None, db if cols else None, dn if nxt else None, R, D, p, 1, 2))
    b = R * D * (2 + 4 + 4 + 4 + (2 if nxt else 0))
    print(f"bwd {name}: {us:.1f} us  {b / us / 1e6:.2f} TB/s")
