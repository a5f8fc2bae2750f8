"""HBM-bound kernel timing probe (not a pytest file): achieved GB/s against algorithmic bytes."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
from b200vit import kernels as K
dev = torch.device("cuda:0")
# usage: perf_elementwise.py [R D]   (default 50432 768: ViT-B/16 at B=256; ViT-L/16-384 at B=64: 36928 1024)
R, D = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) >= 3 else (50432, 768)
M = 4 * D
def timeit(f, n=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
x = torch.randn(R, D, device=dev); gam = torch.ones(D, device=dev); bet = torch.zeros(D, device=dev)
y16 = torch.empty(R, D, device=dev, dtype=torch.bfloat16); mean = torch.empty(R, device=dev); rstd = torch.empty(R, device=dev)
t = timeit(lambda: K.layernorm_fwd(x, D, gam, bet, y16, None, mean, rstd, R, D, 1e-5))
print(f"ln_fwd        {t:7.1f} us  {R*D*6/t/1e3:7.0f} GB/s")
dy = torch.randn(R, D, device=dev).bfloat16(); gin = torch.randn(R, D, device=dev); gout = torch.empty(R, D, device=dev)
g16 = torch.empty(R, D, device=dev, dtype=torch.bfloat16); dg = torch.empty(D, device=dev); db = torch.empty(D, device=dev); dbn = torch.empty(D, device=dev)
t = timeit(lambda: K.layernorm_bwd(dy, None, x, D, mean, rstd, gam, gin, gout, D, None, dg, db, None, R, D))
print(f"ln_bwd        {t:7.1f} us  {R*D*14/t/1e3:7.0f} GB/s")
t = timeit(lambda: K.layernorm_bwd(dy, None, x, D, mean, rstd, gam, gin, gout, D, None, None, None, None, R, D))
print(f"ln_bwd nocols {t:7.1f} us  {R*D*14/t/1e3:7.0f} GB/s")
t = timeit(lambda: K.layernorm_bwd(dy, None, x, D, mean, rstd, gam, gin, gout, D, g16, dg, db, dbn, R, D))
print(f"ln_bwd+next   {t:7.1f} us  {R*D*16/t/1e3:7.0f} GB/s")
t = timeit(lambda: K.layernorm_bwd(dy, None, x, D, mean, rstd, gam, gin, gout, D, g16, dg, db, dbn, R, D, 0.1, 1, 2))
print(f"ln_bwd+next p {t:7.1f} us  {R*D*16/t/1e3:7.0f} GB/s")
da = torch.randn(R, M, device=dev).bfloat16(); o = torch.empty(M, device=dev)
t = timeit(lambda: K.colsum_bf16(da, M, o, R, M))
print(f"colsum[R,M]   {t:7.1f} us  {R*M*2/t/1e3:7.0f} GB/s")
t = timeit(lambda: K.dropout_cast(gin, g16, db, R, D, 0.1, 1, 2))
print(f"dropout_cast  {t:7.1f} us  {R*D*6/t/1e3:7.0f} GB/s")

# calibration: what torch's own elementwise kernels reach on this box
t = timeit(lambda: gout.copy_(gin))
print(f"torch copy f32 {t:7.1f} us  {R*D*8/t/1e3:7.0f} GB/s")
t = timeit(lambda: g16.copy_(gin))
print(f"torch cast     {t:7.1f} us  {R*D*6/t/1e3:7.0f} GB/s")
big = torch.empty(1 << 29, device=dev, dtype=torch.bfloat16); big2 = torch.empty_like(big)
t = timeit(lambda: big2.copy_(big))
print(f"torch copy 1GiB{t:7.1f} us  {big.numel()*4/t/1e3:7.0f} GB/s")
t = timeit(lambda: K.cast_bf16(gin, g16))
print(f"our cast       {t:7.1f} us  {R*D*6/t/1e3:7.0f} GB/s")
