"""Forward-only throughput (eval, inference_mode) with the LayerNorm fold on and off (not a pytest file)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
import b200vit.functional as F_
from b200vit.models import vit, vit_no_classifier
dev = torch.device("cuda:0")
torch.manual_seed(0)
for name, m in (("classifier", vit.ViT(num_classes=1000)), ("headless", vit_no_classifier.ViT())):
    m = m.to(dev).eval()
    for B in (256, 1024):
        X = torch.rand(B, 3, 224, 224, device=dev)
        line = f"{name:10s} B={B:5d}:"
        for flag in (True, False, True, False):
            F_.set_ln_fold(flag)
            with torch.inference_mode():
                for _ in range(3): m(X)
                torch.cuda.synchronize()
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(30): m(X)
                e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 30
            line += f"  fold {'on ' if flag else 'off'} {ms:7.2f} ms ({B / ms * 1e3:8.0f} img/s)"
        print(line, flush=True)
F_.set_ln_fold(True)
