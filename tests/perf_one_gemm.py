"""Launch named GEMMs of the training step a few times, with their real epilogues (for ncu captures and quick timing;
not a pytest file):   python tests/perf_one_gemm.py fc1_fwd fc2_dgrad ...   (names: bench.gemm_shapes_per_step)"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
sys.path.insert(0, str(ROOT))
import torch
import bench
dev = torch.device("cuda:0")
names = sys.argv[1:] or ["fc1_fwd", "fc1_dgrad"]
shapes = {s[0]: s for s in bench.gemm_shapes_per_step(256)}
for n in names:
    name, op, epi, M, N, Kd, _ = shapes[n]
    run = bench.make_gemm_runner(name, op, epi, M, N, Kd, dev)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        run()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 100
    print(f"{n:12s} {us:7.1f} us  {2.0 * M * N * Kd / us / 1e6:7.1f} TFLOP/s")
