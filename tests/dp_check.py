"""Multi-GPU check of b200vit.dp.DataParallel on real GPUs (run with torchrun, not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dp_check.py

Each rank runs the tiny ViT on its shard of a global batch; after backward the (summed) bucket gradients times
1/world must equal the gradient of the same model on the full batch computed by rank 0 alone, and after one
FusedAdam step (grad_scale = 1/world, clip 1.0) all ranks must hold identical weights."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
sys.path.insert(0, str(ROOT / "tests"))

import torch
import torch.distributed as dist

from b200vit.dp import DataParallel
from b200vit.models.vit import ViT
from b200vit.nn import CrossEntropyLoss
from b200vit.optim import FusedAdam
from common import TINY


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(100 + rank)                      # different init per rank: the wrapper must broadcast rank 0's
    model = ViT(**TINY).to(dev)
    dp = DataParallel(model, bucket_mb=0.05)
    g = torch.Generator().manual_seed(5)
    per = 4
    X = torch.rand(per * world, 3, 64, 64, generator=g)
    y = torch.randint(0, 10, (per * world,), generator=g)
    loss_fn = CrossEntropyLoss()
    loss = loss_fn(dp(X[rank * per:(rank + 1) * per].to(dev)), y[rank * per:(rank + 1) * per].to(dev))
    loss.backward()
    grads = {k: p.grad.detach().clone() * dp.grad_scale for k, p in model.named_parameters()}
    # every gradient lives in its bucket; the encoder blocks' 12 x layers gradients were written there directly
    for b in dp.buckets:
        lo, hi = b.flat.data_ptr(), b.flat.data_ptr() + b.flat.numel() * 4
        for p in b.params:
            assert lo <= p.grad.data_ptr() < hi
    ref = ViT(**TINY).to(dev)
    ref.load_state_dict(model.state_dict())
    loss_fn(ref(X.to(dev)), y.to(dev)).backward()
    worst = 0.0
    for k, p in ref.named_parameters():
        if k.endswith("in_proj_bias"):
            continue
        err = ((grads[k] - p.grad).norm() / (p.grad.norm() + 1e-12)).item()
        worst = max(worst, err)
    assert worst < 2e-2, f"rank {rank}: DP gradient differs from full-batch gradient by {worst}"   # bf16 GEMM noise

    opt = FusedAdam(model.parameters(), lr=1e-3, weight_decay=0.03)
    opt.grad_scale = dp.grad_scale
    opt.step(max_norm=1.0)
    flat = torch.cat([p.detach().flatten() for p in model.parameters()])
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    assert all(torch.equal(gathered[0], t) for t in gathered), "weights diverged across ranks after the step"
    # ranks seeded alike still draw different dropout masks (the rank is folded into the mask key)
    from b200vit.models import vit as vitmod
    torch.manual_seed(42)
    pe = vitmod.PatchEmbedding(image_size=64, patch_size=16, embedding_dropout=0.5, embedding_dim=128).to(dev).train()
    img = torch.rand(2, 3, 64, 64, generator=torch.Generator().manual_seed(9)).to(dev)
    torch.manual_seed(42)
    keep = (pe(img) != 0).to(torch.uint8).flatten()
    masks = [torch.empty_like(keep) for _ in range(world)]
    dist.all_gather(masks, keep)
    assert not torch.equal(masks[0], masks[1]), "ranks 0 and 1 drew the same dropout mask"
    # epoch metrics are averaged over ranks
    l, a = dp.reduce_metrics(float(rank), 1.0)
    assert abs(l - (world - 1) / 2) < 1e-12 and a == 1.0
    # gradient accumulation through the engine: 2 micro-batches of 2 == 1 batch of 4 per rank
    from b200vit.going_modular import engine
    xs, ys = X[rank * per:(rank + 1) * per].to(dev), y[rank * per:(rank + 1) * per].to(dev)
    model.zero_grad()
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: 1.0)
    engine.train_step(dp, [(xs[:2], ys[:2]), (xs[2:], ys[2:])], loss_fn, opt, sched, dev, accumulation_steps=2)
    acc_w = torch.cat([p.detach().flatten() for p in model.parameters()])
    gathered = [torch.empty_like(acc_w) for _ in range(world)]
    dist.all_gather(gathered, acc_w)
    assert all(torch.equal(gathered[0], t) for t in gathered), "weights diverged after the accumulated step"
    # the peer-memory all-reduce on its own, against ncclAllReduce on the same values, many calls back to back, with
    # ranks arriving at different times
    expect = os.environ.get("B200VIT_DP_ALLREDUCE", "peer")
    assert (dp.peer is not None) == (expect == "peer" and world <= 8), f"transport: peer={dp.peer is not None}, asked {expect}"
    if dp.peer is not None:
        gen = torch.Generator(device=dev).manual_seed(1000 + rank)
        for it in range(12):
            b = dp.buckets[it % len(dp.buckets)]
            b.flat.copy_(torch.randn(b.flat.numel(), device=dev, generator=gen))
            want = b.flat.clone()
            dist.all_reduce(want)
            if it % 3 == rank % 3:
                torch.cuda._sleep(3_000_000)          # this rank's bucket is late
            dp.peer.all_reduce(b.flat, b.ready, b.done).wait()
            err = ((b.flat - want).abs().max() / want.abs().max()).item()
            assert err < 1e-6, f"rank {rank}: peer all-reduce differs from NCCL by {err} (call {it})"
            if world == 2:
                assert torch.equal(b.flat, want)
            got = [torch.empty_like(b.flat) for _ in range(world)]
            dist.all_gather(got, b.flat)
            assert all(torch.equal(got[0], t) for t in got), "ranks hold different sums"
    # LAST (it leaves the gradients in an undefined state): a second synchronised backward without zero_grad() must
    # be refused -- it would reduce the first sum again
    model.zero_grad()
    loss_fn(dp(xs), ys).backward()
    try:
        loss_fn(dp(xs), ys).backward()
        raise AssertionError("double synchronised backward was not detected")
    except RuntimeError as e:
        assert "already all-reduced" in str(e), e
    torch.cuda.synchronize()
    if rank == 0:
        print(f"dp_check ok: world={world} buckets={len(dp.buckets)} transport={'peer' if dp.peer is not None else 'nccl'} "
              f"worst grad rel err {worst:.2e}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
