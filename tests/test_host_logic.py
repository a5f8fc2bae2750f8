"""CPU tests of the host-side logic that does not depend on the CUDA library: the engine loop (with a stock torch
model / optimizer, as the reference engine allows any nn.Module) and the data-parallel wrapper over gloo with
world_size 2."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_engine_loop_matches_reference_semantics_on_cpu():
    """Mean of per-batch means, per-batch LR stepping, clip at 1.0: checked against a hand-rolled copy of the
    reference loop (going_modular/engine.py:45-79) on a small torch model."""
    from b200vit.going_modular import engine
    torch.manual_seed(0)
    g = torch.Generator().manual_seed(1)
    data = [(torch.randn(8, 6, generator=g), torch.randint(0, 3, (8,), generator=g)) for _ in range(4)]
    data.append((torch.randn(3, 6, generator=g), torch.randint(0, 3, (3,), generator=g)))   # ragged last batch

    def make():
        torch.manual_seed(2)
        m = nn.Sequential(nn.Linear(6, 16), nn.Tanh(), nn.Linear(16, 3))
        opt = torch.optim.Adam(m.parameters(), lr=1e-2, weight_decay=0.03)
        sched = torch.optim.lr_scheduler.LinearLR(opt, start_factor=0.1, total_iters=5)
        return m, opt, sched

    m, opt, sched = make()
    loss_fn = nn.CrossEntropyLoss()
    tl, ta = engine.train_step(m, data, loss_fn, opt, sched, torch.device("cpu"))

    r, ropt, rsched = make()
    rl = ra = 0.0
    r.train()
    for X, y in data:
        pred = r(X)
        loss = loss_fn(pred, y)
        rl += loss.item()
        ropt.zero_grad()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(r.parameters(), max_norm=1.0)
        ropt.step()
        rsched.step()
        ra += (torch.argmax(torch.softmax(pred, dim=1), dim=1) == y).sum().item() / len(pred)
    assert tl == pytest.approx(rl / len(data), rel=1e-6)
    assert ta == pytest.approx(ra / len(data), rel=1e-6)
    for a, b in zip(m.parameters(), r.parameters()):
        assert torch.allclose(a, b, atol=1e-7)
    el, ea = engine.test_step(m, data, loss_fn, torch.device("cpu"))
    with torch.inference_mode():
        rl = sum(loss_fn(r(X), y).item() for X, y in data) / len(data)
    assert el == pytest.approx(rl, rel=1e-6)
    res = engine.train(m, data[:2], data[:1], opt, loss_fn, sched, epochs=2, device=torch.device("cpu"))
    assert set(res) == {"train_loss", "train_acc", "test_loss", "test_acc"} and all(len(v) == 2 for v in res.values())


def _dp_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from b200vit.dp import DataParallel
        torch.manual_seed(100 + rank)                      # different init per rank: broadcast must fix it
        net = nn.Sequential(nn.Linear(10, 32), nn.ReLU(), nn.Linear(32, 32), nn.ReLU(), nn.Linear(32, 4))
        dp = DataParallel(net, bucket_mb=0.002)            # tiny buckets -> several of them
        assert len(dp.buckets) >= 2
        g = torch.Generator().manual_seed(7)
        X = torch.randn(8 * world, 10, generator=g)
        y = torch.randint(0, 4, (8 * world,), generator=g)
        xs, ys = X[rank * 8:(rank + 1) * 8], y[rank * 8:(rank + 1) * 8]
        loss = nn.functional.cross_entropy(dp(xs), ys)
        loss.backward()
        dp.average_gradients()
        # single-process reference on the full batch with rank-0 weights
        ref = nn.Sequential(nn.Linear(10, 32), nn.ReLU(), nn.Linear(32, 32), nn.ReLU(), nn.Linear(32, 4))
        ref.load_state_dict(net.state_dict())
        nn.functional.cross_entropy(ref(X), y).backward()
        err = max((a.grad - b.grad).abs().max().item() for a, b in zip(net.parameters(), ref.parameters()))
        sd = torch.cat([p.detach().flatten() for p in net.parameters()])
        gathered = [torch.zeros_like(sd) for _ in range(world)]
        dist.all_gather(gathered, sd)
        same = all(torch.equal(gathered[0], t) for t in gathered)
        # second step with zero_grad(set_to_none=True) and with no_sync accumulation
        for p in net.parameters():
            p.grad = None
        with dp.no_sync():
            nn.functional.cross_entropy(dp(xs[:4]), ys[:4], reduction="sum").backward()
        nn.functional.cross_entropy(dp(xs[4:]), ys[4:], reduction="sum").backward()
        ref.zero_grad()
        nn.functional.cross_entropy(ref(X), y, reduction="sum").backward()
        err2 = max((a.grad - b.grad).abs().max().item() for a, b in zip(net.parameters(), ref.parameters()))
        out.put((rank, err, same, err2))
    finally:
        dist.destroy_process_group()


def test_data_parallel_gloo_world2():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dp_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, same, err2 in res:
        assert same, "parameters differ across ranks after broadcast"
        assert err < 1e-6, f"rank {rank}: averaged DP gradient differs from full-batch gradient by {err}"
        assert err2 < 1e-5, f"rank {rank}: accumulated (no_sync) DP gradient differs by {err2}"
