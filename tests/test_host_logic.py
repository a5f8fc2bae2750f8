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


class _FakeArenaTransport:
    """Stands in for dp._PeerTransport on CPU: one arena for all buckets (as the symmetric allocation is), the sum done
    by gloo. Exercises the arena slicing / padding / view logic of DataParallel without a GPU."""

    def __init__(self, floats):
        self.arena = torch.zeros(floats)
        self.calls = 0

    def make_events(self):
        return None, None

    def all_reduce(self, flat, ready, done):
        assert flat.numel() % 64 == 0 and flat.data_ptr() >= self.arena.data_ptr()
        assert flat.data_ptr() + flat.numel() * 4 <= self.arena.data_ptr() + self.arena.numel() * 4
        self.calls += 1
        return dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=True)


def _dp_worker(rank, world, port, out, arena=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from b200vit import dp as dpmod
        from b200vit.dp import DataParallel
        if arena:
            def fake(self, groups):
                return _FakeArenaTransport(sum(dpmod._padded(sum(p.numel() for p in g)) for g in groups))
            DataParallel._make_peer_transport = fake
        torch.manual_seed(100 + rank)                      # different init per rank: broadcast must fix it
        net = nn.Sequential(nn.Linear(10, 32), nn.ReLU(), nn.Linear(32, 32), nn.ReLU(), nn.Linear(32, 4))
        dp = DataParallel(net, bucket_mb=0.002)            # tiny buckets -> several of them
        assert len(dp.buckets) >= 2
        if arena:       # every bucket is a 256-byte aligned, zero-padded slice of the one arena, back to back
            off = 0
            for b in dp.buckets:
                n = sum(p.numel() for p in b.params)
                assert b.flat.data_ptr() == dp.peer.arena.data_ptr() + 4 * off and b.flat.numel() == dpmod._padded(n)
                off += b.flat.numel()
            assert off == dp.peer.arena.numel()
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
        # a second synchronised backward on top of reduced gradients is refused
        refused = False
        try:
            nn.functional.cross_entropy(dp(xs), ys).backward()
        except RuntimeError as e:
            refused = "already all-reduced" in str(e)
        assert refused
        for b in dp.buckets:
            if b.work is not None:
                b.work.wait()
                b.work = None
            b.pending = len(b.params)
        dp._callback_queued = False
        l, a = dp.reduce_metrics(float(rank + 1), 0.5)
        assert abs(l - (world + 1) / 2) < 1e-12 and a == 0.5
        from b200vit import functional as F_
        assert F_._dropout_stream == rank
        if arena:
            assert dp.peer.calls >= 2 * len(dp.buckets)
            for b in dp.buckets:            # the padding behind each bucket's parameters stays zero
                n = sum(p.numel() for p in b.params)
                assert float(b.flat[n:].abs().sum()) == 0.0
        out.put((rank, err, same, err2))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("arena", [False, True])
def test_data_parallel_gloo_world2(arena):
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dp_worker, args=(r, world, port, out, arena)) for r in range(world)]
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


class _StubLib:
    """Test double for libb200vit.so: every entry point succeeds without doing anything (host logic only)."""

    def __getattr__(self, name):
        if name == "b200vit_last_error":
            return lambda: b""
        return lambda *a: 0


def test_fused_adam_step_counter_survives_state_dict_round_trip():
    """All parameters of a FusedAdam share ONE step tensor; state_dict()/load_state_dict() keep that aliasing. The
    step must advance by one per optimizer step before and after a resume (not once per parameter)."""
    from b200vit import _C
    from b200vit.optim import FusedAdam
    _C.set_library(_StubLib())
    try:
        def make():
            torch.manual_seed(0)
            return [nn.Parameter(torch.randn(4, 3)) for _ in range(5)]

        def step(ps, opt, n):
            for _ in range(n):
                for p in ps:
                    p.grad = torch.ones_like(p)
                opt.step(max_norm=1.0)

        ps = make()
        opt = FusedAdam(ps, lr=1e-3)
        step(ps, opt, 3)
        assert {int(opt.state[p]["step"]) for p in ps} == {3}
        import io
        buf = io.BytesIO()
        torch.save(opt.state_dict(), buf)
        buf.seek(0)
        ps2 = make()
        opt2 = FusedAdam(ps2, lr=1e-3)
        opt2.load_state_dict(torch.load(buf))
        assert {int(opt2.state[p]["step"]) for p in ps2} == {3}
        step(ps2, opt2, 3)
        assert {int(opt2.state[p]["step"]) for p in ps2} == {6}
        # a parameter that joins after the first step counts from its own zero, as in torch.optim.Adam
        late = nn.Parameter(torch.randn(2))
        opt2.add_param_group({"params": [late]})
        late.grad = torch.ones_like(late)
        step(ps2, opt2, 1)
        assert int(opt2.state[late]["step"]) == 1 and {int(opt2.state[p]["step"]) for p in ps2} == {7}
    finally:
        _C.set_library(None)


def test_engine_gradient_accumulation_equals_one_large_batch():
    """train_step(accumulation_steps=k): k micro-batches -> one optimizer step with the averaged gradient."""
    from b200vit.going_modular import engine
    g = torch.Generator().manual_seed(3)
    X, y = torch.randn(16, 6, generator=g), torch.randint(0, 3, (16,), generator=g)

    def make():
        torch.manual_seed(2)
        m = nn.Sequential(nn.Linear(6, 16), nn.Tanh(), nn.Linear(16, 3))
        opt = torch.optim.SGD(m.parameters(), lr=0.1)
        return m, opt, torch.optim.lr_scheduler.LinearLR(opt, start_factor=0.5, total_iters=4)

    m1, o1, s1 = make()
    engine.train_step(m1, [(X[:8], y[:8]), (X[8:], y[8:])], nn.CrossEntropyLoss(), o1, s1, torch.device("cpu"))
    ma, oa, sa = make()
    micro = [(X[i:i + 4], y[i:i + 4]) for i in range(0, 16, 4)]
    la, _ = engine.train_step(ma, micro, nn.CrossEntropyLoss(), oa, sa, torch.device("cpu"), accumulation_steps=2)
    for a, b in zip(m1.parameters(), ma.parameters()):
        assert torch.allclose(a, b, atol=1e-6)
    assert sa.last_epoch == s1.last_epoch == 2                  # the schedule advances per optimizer step
    with pytest.raises(ValueError):
        engine.train_step(ma, micro, nn.CrossEntropyLoss(), oa, sa, torch.device("cpu"), accumulation_steps=0)


def test_dropout_keys_follow_seed_and_stream():
    """Dropout keys: reproducible from torch.manual_seed, fresh per call, different per data-parallel rank."""
    from b200vit import functional as F_
    torch.manual_seed(7)
    a = [F_.next_dropout_key() for _ in range(3)]
    torch.manual_seed(7)
    b = [F_.next_dropout_key() for _ in range(3)]
    assert a == b and len(set(a)) == 3
    prev = F_.set_dropout_stream(1)
    try:
        torch.manual_seed(7)
        c = [F_.next_dropout_key() for _ in range(3)]
    finally:
        F_.set_dropout_stream(prev)
    assert all(x[0] != y[0] and x[1] == y[1] for x, y in zip(a, c))
