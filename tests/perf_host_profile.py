"""cProfile of the host side of training steps at a small batch (not a pytest file): where the Python time goes."""
import cProfile, pstats, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
from b200vit.models.vit import ViT
from b200vit.nn import CrossEntropyLoss
from b200vit.optim import FusedAdam
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = torch.device("cuda:0")
torch.manual_seed(0)
m = ViT(num_classes=3).to(dev).train()
opt = FusedAdam(m.parameters(), lr=1e-4)
lf = CrossEntropyLoss()
X = torch.rand(B, 3, 224, 224, device=dev); y = torch.randint(0, 3, (B,), device=dev)
def step():
    loss = lf(m(X), y); opt.zero_grad(); loss.backward(); opt.step(max_norm=1.0)
for _ in range(5): step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(20): step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr); st.sort_stats("cumulative").print_stats(28)
