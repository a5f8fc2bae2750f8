"""In-situ kernel timeline of a few training steps via torch.profiler (CUPTI): busy time vs wall time and the
largest idle gaps (not a pytest file)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
from torch.profiler import ProfilerActivity, profile
from b200vit.models.vit import ViT
from b200vit.nn import CrossEntropyLoss
from b200vit.optim import FusedAdam

# usage: perf_timeline.py [vitl]   (default: ViT-B/16 3-class B=256; vitl: ViT-L/16 at 384x384, B=64)
VITL = len(sys.argv) > 1 and sys.argv[1] == "vitl"
B, IMG, NC = (64, 384, 1000) if VITL else (256, 224, 3)
dev = torch.device("cuda:0")
torch.manual_seed(0)
kw = dict(image_size=384, num_transformer_layer=24, embedding_dim=1024, num_heads=16, mlp_size=4096) if VITL else {}
m = ViT(num_classes=NC, **kw).to(dev).train()
opt = FusedAdam(m.parameters(), lr=1e-4)
lf = CrossEntropyLoss()
X = torch.rand(B, 3, IMG, IMG, device=dev); y = torch.randint(0, NC, (B,), device=dev)
def step():
    loss = lf(m(X), y); opt.zero_grad(); loss.backward(); opt.step(max_norm=1.0)
for _ in range(4): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ks = sorted([(e.time_range.start, e.time_range.end, e.name) for e in evs])
busy = sum(e - s for s, e, _ in ks)
wall = ks[-1][1] - ks[0][0]
print(f"kernels {len(ks)}  busy {busy/3e3:.2f} ms/step  wall {wall/3e3:.2f} ms/step  idle {(wall-busy)/3e3:.2f} ms/step")
gaps = sorted([(ks[i+1][0] - ks[i][1], ks[i][2][:50], ks[i+1][2][:50]) for i in range(len(ks)-1)], reverse=True)
print("gap histogram (us): >100:", sum(g[0] > 100 for g in gaps), " 20-100:", sum(20 < g[0] <= 100 for g in gaps),
      " 5-20:", sum(5 < g[0] <= 20 for g in gaps), " <=5:", sum(g[0] <= 5 for g in gaps))
print("sum of gaps >20us: %.2f ms, 5-20us: %.2f ms, <=5us: %.2f ms (over 3 steps)" % (
    sum(g[0] for g in gaps if g[0] > 20) / 1e3, sum(g[0] for g in gaps if 5 < g[0] <= 20) / 1e3,
    sum(g[0] for g in gaps if g[0] <= 5) / 1e3))
for g in gaps[:12]:
    print(f"{g[0]:8.1f} us  after {g[1]}  before {g[2]}")
import collections
agg = collections.defaultdict(float)
for s, e, n in ks: agg[n[:60]] += (e - s)
for n, t in sorted(agg.items(), key=lambda kv: -kv[1])[:24]:
    print(f"{t/3e3:8.3f} ms/step  {n}")
