"""Host-overhead probe (not a pytest file): CPU enqueue time vs GPU time of one training step."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
from b200vit.models.vit import ViT
from b200vit.nn import CrossEntropyLoss
from b200vit.optim import FusedAdam

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = ViT(num_classes=3).to(dev).train()
opt = FusedAdam(m.parameters(), lr=1e-4)
lf = CrossEntropyLoss()
X = torch.rand(B, 3, 224, 224, device=dev); y = torch.randint(0, 3, (B,), device=dev)

def step():
    loss = lf(m(X), y)
    opt.zero_grad()
    loss.backward()
    opt.step(max_norm=1.0)

for _ in range(3): step()
torch.cuda.synchronize()
n = 10
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for _ in range(n): step()
t1 = time.perf_counter(); e1.record(); torch.cuda.synchronize()
print(f"B={B}: cpu enqueue {1e3*(t1-t0)/n:.2f} ms/step, gpu {e0.elapsed_time(e1)/n:.2f} ms/step")
# forward-only / backward split
torch.cuda.synchronize(); e0.record()
for _ in range(n):
    with torch.no_grad(): m(X)
e1.record(); torch.cuda.synchronize()
print(f"forward-only (no_grad, train-mode dropout) {e0.elapsed_time(e1)/n:.2f} ms")
