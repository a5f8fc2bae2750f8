"""Attention kernel timing probe (not a pytest file). B200VIT_ATTN_MODE=1 selects the mma.sync kernels."""
import math, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
from b200vit import kernels as K
dev = torch.device("cuda:0")
# usage: perf_attn.py [B T H]   (default: the ViT-B/16 case 256 197 12; ViT-L/16-384: 64 577 16)
B, T, H = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (256, 197, 12)
hd = 64
qkv = torch.randn(B, T, 3, H, hd, device=dev).bfloat16()
out = torch.empty(B, T, H * hd, device=dev, dtype=torch.bfloat16)
lse = torch.empty(B, H, T, device=dev)
dout = torch.randn(B, T, H * hd, device=dev).bfloat16()
delta = torch.empty(B, H, T, device=dev)
dqkv = torch.empty_like(qkv)
sc = 1 / math.sqrt(hd)
def timeit(f, n=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
f = timeit(lambda: K.attention_fwd(qkv, out, lse, B, T, H, hd, sc))
b = timeit(lambda: K.attention_bwd(qkv, out, dout, lse, delta, dqkv, B, T, H, hd, sc))
dbias = torch.empty(3 * H * hd, device=dev)
bb = timeit(lambda: K.attention_bwd(qkv, out, dout, lse, delta, dqkv, B, T, H, hd, sc, dbias=dbias))
fl = 4 * B * H * T * T * hd
print(f"attention fwd {f:.1f} us ({fl/f/1e6:.0f} TFLOP/s)  bwd {b:.1f} us ({2.5*fl/b/1e6:.0f} TFLOP/s algorithmic)  bwd+bias-grad {bb:.1f} us")
