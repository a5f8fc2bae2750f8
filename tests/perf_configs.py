"""Throughput probe for BASELINE.json configs 3-5 plus an eager-PyTorch GPU baseline of config 2 (not a pytest file).

  C3  ViT-B/16, 1000 classes, per-GPU batch 512 (one GPU's share of global 4096)
  C4  ViT-L/16 at 384x384 (577 tokens), per-GPU batch 64
  C5  headless ViT-B/16 forward sweep, batch 1..1024, inference_mode
  E   the same architecture written with stock torch.nn modules (nn.MultiheadAttention, nn.GELU, ...), B=256,
      one training step in fp32 (TF32 off) and under bf16 autocast -- "eager PyTorch on the same B200".
Prints one JSON object per line. Usage: python tests/perf_configs.py [c3] [c4] [c5] [eager]
"""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch
from torch import nn

from b200vit.models.vit import ViT
from b200vit.models import vit_no_classifier
from b200vit.nn import CrossEntropyLoss
from b200vit.optim import FusedAdam

dev = torch.device("cuda:0")
which = set(sys.argv[1:]) or {"c3", "c4", "c5", "eager"}


def train_flops_per_image(T, D, L, M, C, patch_k):
    """2mnk per GEMM, attention 2*H*T^2*hd for QK^T and PV; backward = 2x (patch: 1x)."""
    patch = 2.0 * (T - 1) * patch_k * D
    layer = 2.0 * T * D * 3 * D + 2 * 2.0 * T * T * D + 2.0 * T * D * D + 2 * 2.0 * T * D * M
    head = 2.0 * D * C
    return patch * 2 + 3 * (L * layer + head), patch + L * layer + head


def timed(fn, warm, n):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def train_probe(name, kwargs, B, img, flops):
    torch.manual_seed(0)
    m = ViT(**kwargs).to(dev).train()
    opt = FusedAdam(m.parameters(), lr=1e-4)
    lf = CrossEntropyLoss()
    X = torch.rand(B, 3, img, img, device=dev)
    y = torch.randint(0, kwargs["num_classes"], (B,), device=dev)

    def step():
        loss = lf(m(X), y)
        opt.zero_grad()
        loss.backward()
        opt.step(max_norm=1.0)

    ms = timed(step, 3, 8)
    ips = B / ms * 1e3
    print(json.dumps({"config": name, "batch": B, "ms_per_step": round(ms, 2), "images_per_s": round(ips, 1),
                      "tflops": round(ips * flops / 1e12, 1), "mfu_vs_2250": round(ips * flops / 2250e12, 4),
                      "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 1)}), flush=True)
    del m, opt, X, y
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()


if "c3" in which:
    f, _ = train_flops_per_image(197, 768, 12, 3072, 1000, 768)
    train_probe("C3 ViT-B/16 1000-class, per-GPU batch 512", dict(num_classes=1000), 512, 224, f)

if "c4" in which:
    f, _ = train_flops_per_image(577, 1024, 24, 4096, 1000, 768)
    train_probe("C4 ViT-L/16 384x384 (577 tokens), per-GPU batch 64",
                dict(image_size=384, num_transformer_layer=24, embedding_dim=1024, num_heads=16, mlp_size=4096,
                     num_classes=1000), 64, 384, f)

if "c5" in which:
    torch.manual_seed(0)
    m = vit_no_classifier.ViT().to(dev).eval()
    from b200vit.inference import GraphedForward
    graphed = GraphedForward(m)
    _, ffwd = train_flops_per_image(197, 768, 12, 3072, 0, 768)
    with torch.inference_mode():
        for B in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024):
            X = torch.rand(B, 3, 224, 224, device=dev)
            ms = timed(lambda: m(X), 3, 10 if B <= 256 else 5)
            ips = B / ms * 1e3
            rec = {"config": "C5 headless ViT-B/16 forward", "batch": B, "ms": round(ms, 3),
                   "images_per_s": round(ips, 1), "tflops": round(ips * ffwd / 1e12, 1)}
            if B <= 64:         # launch-bound range: the same forward replayed from a CUDA graph
                msg = timed(lambda: graphed(X), 3, 20)
                rec.update({"graphed_ms": round(msg, 3), "graphed_images_per_s": round(B / msg * 1e3, 1)})
            print(json.dumps(rec), flush=True)
    del m
    torch.cuda.empty_cache()

if "eager" in which:
    class Block(nn.Module):
        def __init__(self, D=768, H=12, M=3072, p=0.1):
            super().__init__()
            self.ln1, self.ln2 = nn.LayerNorm(D), nn.LayerNorm(D)
            self.attn = nn.MultiheadAttention(D, H, dropout=0.0, batch_first=True)
            self.mlp = nn.Sequential(nn.Linear(D, M), nn.GELU(), nn.Dropout(p), nn.Linear(M, D), nn.Dropout(p))

        def forward(self, x):
            h = self.ln1(x)
            x = self.attn(h, h, h, need_weights=False)[0] + x
            return self.mlp(self.ln2(x)) + x

    class EagerViT(nn.Module):
        def __init__(self, D=768, L=12, C=3):
            super().__init__()
            self.conv = nn.Conv2d(3, D, 16, 16)
            self.cls = nn.Parameter(torch.rand(1, 1, D))
            self.pos = nn.Parameter(torch.rand(1, 197, D))
            self.drop = nn.Dropout(0.1)
            self.blocks = nn.Sequential(*[Block() for _ in range(L)])
            self.ln = nn.LayerNorm(D)
            self.head = nn.Linear(D, C)

        def forward(self, x):
            x = self.conv(x).flatten(2).permute(0, 2, 1)
            x = torch.cat((self.cls.expand(x.shape[0], -1, -1), x), dim=1) + self.pos
            x = self.blocks(self.drop(x))
            return self.head(self.ln(x)[:, 0])

    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    f, _ = train_flops_per_image(197, 768, 12, 3072, 3, 768)
    for mode in ("fp32", "bf16_autocast"):
        torch.manual_seed(0)
        m = EagerViT().to(dev).train()
        opt = torch.optim.Adam(m.parameters(), lr=1e-4)
        lf = nn.CrossEntropyLoss()
        B = 256
        X = torch.rand(B, 3, 224, 224, device=dev)
        y = torch.randint(0, 3, (B,), device=dev)

        def step():
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=mode != "fp32"):
                loss = lf(m(X), y)
            opt.zero_grad()
            loss.backward()
            torch.nn.utils.clip_grad_norm_(m.parameters(), 1.0)
            opt.step()

        ms = timed(step, 3, 6)
        ips = B / ms * 1e3
        print(json.dumps({"config": f"E eager torch.nn ViT-B/16 3-class, {mode}", "batch": B, "ms_per_step": round(ms, 2),
                          "images_per_s": round(ips, 1), "tflops": round(ips * f / 1e12, 1),
                          "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 1)}), flush=True)
        del m, opt, X, y
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
