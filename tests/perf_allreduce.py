"""All-reduce of gradient-bucket-sized fp32 buffers alone on the GPUs (no overlapping compute): ncclAllReduce against
the library's peer-memory kernel, CUDA events, 10 back-to-back calls (not a pytest file).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29516 tests/perf_allreduce.py
"""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
from b200vit.dp import _PeerTransport  # noqa: E402

local = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
world, rank = dist.get_world_size(), dist.get_rank()
sizes = (28, 100, 343)
peer = _PeerTransport(int(max(sizes) * 1e6 / 4) // 64 * 64, dev, None, world, rank)


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for mb in sizes:
    n = int(mb * 1e6 / 4) // 64 * 64
    t = torch.ones(n, device=dev)
    us = timed(lambda: dist.all_reduce(t))
    if rank == 0:
        print(f"nccl all_reduce {mb} MB fp32: {us:.0f} us  ({mb / us * 1e3:.0f} GB/s algbw)", flush=True)
    flat = peer.arena[:n]
    ready, done = torch.cuda.Event(), torch.cuda.Event()
    for ctas in (74, 148, 296, 592):
        peer.ctas = ctas
        us = timed(lambda: peer.all_reduce(flat, ready, done).wait())
        if rank == 0:
            print(f"peer all_reduce {mb} MB fp32, {ctas} CTAs: {us:.0f} us  ({mb / us * 1e3:.0f} GB/s algbw)", flush=True)
dist.destroy_process_group()
