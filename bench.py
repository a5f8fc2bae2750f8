#!/usr/bin/env python
"""Headline benchmark: ViT-B/16 training images/sec on B200 (BASELINE.json metric), one JSON line on rank 0.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...     # the unmodified reference's engine.train_step on the host cores

Workload at N=1: BASELINE.json configs[1] -- ViT-B/16, 3-class head, bf16 GEMM operands / fp32 accumulation,
batch 256, synthetic 224x224 images, dropout as shipped (0.1 / 0.1 / 0), the notebook's optimizer recipe
(Adam with coupled L2 0.03 on matrices, warm-up + linear decay stepped per batch, clip at global norm 1) run through
b200vit.going_modular.engine.train_step. N>1: the same per-GPU batch on every rank (weak scaling) under
b200vit.dp.DataParallel, gradients summed over NVLink peer memory by the library's own all-reduce kernel (NCCL as
fallback) and averaged inside FusedAdam; `data_parallel` in the line also times the same steps without the exchange,
per rank, so the weak-scaling loss splits into "slowest GPU of the box" and "exchange".

A "step" = one optimizer step = one batch through forward, loss, backward, clip, Adam, LR schedule.
  value : images/sec with the batches already resident in HBM (CUDA-event timed, max over ranks)
  e2e   : the same metric through the public API with HOST (pinned) batches: the per-step H2D copy and the
          per-step D2H read of loss/accuracy are inside the timed region
  roofline : the tcgen05 GEMM kernel (95 % of the FLOPs): algorithmic FLOPs of every GEMM launch of one step /
             CUDA-event time of those very launches INSIDE real training steps (events on the launching stream around
             every C-ABI call of two extra steps), vs the measured SUSTAINED bf16 peak; `isolated` repeats it with each
             shape replayed back to back on its own, vs the measured BURST peak; `per_kernel` gives attention,
             LayerNorm and Adam against their own bounds
  cpu_baseline : the UNMODIFIED reference engine.train_step + models/vit.py (baseline/_ref, kind "reference"; the oracle
             port only if that copy is missing) timed on the host cores, bounded sample
  gpu_eager_baseline : that same unmodified reference on this B200 in eager PyTorch, fp32 (TF32 off) and bf16 autocast
  config3 / config4 : BASELINE.json configs[2] (1000 classes, global batch 4096 = N x k x 512 with gradient
             accumulation under no_sync) at every N, and configs[3] (ViT-L/16 at 384 px, 64 images per GPU) at N = 8
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
for _p in (str(ROOT / "pytorch-vit-paper-replication_b200"), str(ROOT)):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

METRIC = "ViT-B/16 train images/sec"
UNIT = "images/sec"
NOMINAL_BF16_TFLOPS = 2250.0


# ---------------------------------------------------------------------------------------------------------
# FLOP model: 2mnk per GEMM, attention 2*H*T^2*hd per QK^T and per PV, backward = 2x forward
# except the patch embedding (no image gradient); elementwise work not counted.
# ---------------------------------------------------------------------------------------------------------
def train_gflop_per_image(image=224, patch=16, layers=12, d=768, heads=12, mlp=3072, classes=3):
    t = (image // patch) ** 2 + 1
    npatch = t - 1
    hd = d // heads
    patch_f = 2 * npatch * d * (3 * patch * patch)
    layer = 2 * t * d * 3 * d + 2 * 2 * heads * t * t * hd + 2 * t * d * d + 2 * 2 * t * d * mlp
    head = 2 * d * classes
    fwd = patch_f + layers * layer + head
    train = 2 * patch_f + 3 * (layers * layer + head)
    return fwd / 1e9, train / 1e9


def gemm_shapes_per_step(batch, image=224, patch=16, layers=12, d=768, mlp=3072):
    """(op, epi, M, N, K, count) of every tcgen05 GEMM launch in one training step."""
    from b200vit import kernels as K
    t = (image // patch) ** 2 + 1
    r = batch * t
    rp = batch * (t - 1)
    kp = 3 * patch * patch
    L = layers
    return [
        ("patch_fwd", K.OP_NT, K.EPI_PATCH, rp, d, kp, 1), ("patch_wgrad", K.OP_TN, K.EPI_ACCUM, d, kp, rp, 1),
        ("qkv_fwd", K.OP_NT, K.EPI_BF16, r, 3 * d, d, L), ("proj_fwd", K.OP_NT, K.EPI_BIAS_RES, r, d, d, L),
        ("fc1_fwd", K.OP_NT, K.EPI_BIAS_GELU, r, mlp, d, L), ("fc2_fwd", K.OP_NT, K.EPI_BIAS_RES, r, d, mlp, L),
        ("fc2_wgrad", K.OP_TN, K.EPI_ACCUM, d, mlp, r, L), ("fc2_dgrad", K.OP_NN, K.EPI_DGELU, r, mlp, d, L),
        ("fc1_wgrad", K.OP_TN, K.EPI_ACCUM, mlp, d, r, L), ("fc1_dgrad", K.OP_NN, K.EPI_BF16, r, d, mlp, L),
        ("proj_wgrad", K.OP_TN, K.EPI_ACCUM, d, d, r, L), ("proj_dgrad", K.OP_NN, K.EPI_BF16, r, d, d, L),
        ("qkv_wgrad", K.OP_TN, K.EPI_ACCUM, 3 * d, d, r, L), ("qkv_dgrad", K.OP_NN, K.EPI_BF16, r, d, 3 * d, L),
    ]


def make_gemm_runner(name, op, epi, M, N, Kd, dev):
    """Operands + a closure that launches one GEMM of the training step with its real epilogue (dropout 0.1 where
    the step has it)."""
    from b200vit import kernels as K
    if op == K.OP_NT:
        A = torch.randn(M, Kd, device=dev).bfloat16(); B = torch.randn(N, Kd, device=dev).bfloat16(); lda, ldb = Kd, Kd
    elif op == K.OP_NN:
        A = torch.randn(M, Kd, device=dev).bfloat16(); B = torch.randn(Kd, N, device=dev).bfloat16(); lda, ldb = Kd, N
    else:
        A = torch.randn(Kd, M, device=dev).bfloat16(); B = torch.randn(Kd, N, device=dev).bfloat16(); lda, ldb = M, N
    f32 = epi in (K.EPI_BIAS_RES, K.EPI_ACCUM, K.EPI_PATCH)
    rows_out = M + M // 196 + 1 if epi == K.EPI_PATCH else M
    C = torch.zeros(rows_out, N, device=dev, dtype=torch.float32 if f32 else torch.bfloat16)
    C2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16) if epi == K.EPI_BIAS_GELU else None
    bias = torch.randn(N, device=dev) if op == K.OP_NT else None
    aux, ldaux, period, p = None, 0, 0, 0.0
    if epi == K.EPI_BIAS_RES:
        aux, ldaux = torch.randn(M, N, device=dev), N
    if epi == K.EPI_DGELU:
        aux, ldaux = torch.randn(M, N, device=dev).bfloat16(), N
    if epi == K.EPI_PATCH:
        aux, ldaux, period = torch.randn(197, N, device=dev), N, 196
    if name in ("fc1_fwd", "fc2_fwd", "fc2_dgrad", "patch_fwd"):
        p = 0.1
    split = 0 if epi == K.EPI_ACCUM else 1

    def run():
        K.gemm(op, epi, M, N, Kd, A, lda, B, ldb, C, N, C2=C2, ldc2=N, bias=bias, aux=aux, ldaux=ldaux,
               dropout_p=p, seed=1, offset=2, row_period=period, split_k=split)
    return run


def gemm_traffic_per_launch(batch):
    """Count-weighted mean DRAM bytes (read + write) per GEMM launch of one step, from the committed ncu --set full
    capture (profiles/r02_gemm_traffic.json, written by tools/summarize_ncu.py); None when the file is absent or
    was taken at another batch size."""
    path = ROOT / "profiles" / "r02_gemm_traffic.json"
    if not path.exists():
        path = ROOT / "profiles" / "r01_gemm_traffic.json"
    if not path.exists():
        return None, None
    rec = json.loads(path.read_text())
    if rec.get("batch") != batch:
        return None, None
    tot, n = 0.0, 0
    for name, *_rest, count in gemm_shapes_per_step(batch):
        if name not in rec["dram_bytes"]:
            return None, None
        tot += rec["dram_bytes"][name] * count
        n += count
    return tot / n, rec.get("source")


def measure_step_kernels(run_steps, n_steps, batch, peaks, classes):
    """CUDA events around every C-ABI call of `n_steps` real training steps -> in-step time per kernel family, the
    GEMM roofline (FLOP-weighted over the step's 146 launches) and the other kernels against their own bounds."""
    from b200vit import _C
    from b200vit import functional as F_
    prev = F_.set_composite(False)      # one C call per kernel, so that every launch gets its own pair of events
    try:
        with _C.CallTimer() as tm:
            run_steps(n_steps)
    finally:
        F_.set_composite(prev)
    times = tm.by_name()
    by = {k: sum(v) / n_steps for k, v in times.items()}          # ms per step
    calls = {k: len(v) / n_steps for k, v in times.items()}
    g = tm.gemms()
    gemm_ms = sum(x[-1] for x in g) / n_steps
    gemm_flops = sum(2.0 * x[2] * x[3] * x[4] for x in g) / n_steps
    n_launch = len(g) / n_steps
    sustained = peaks.get("bf16_tflops_sustained") or 1400.0
    hbm = peaks.get("hbm_gbs") or 6500.0
    T, D, H, L = 197, 768, 12, 12
    R = batch * T
    attn_fwd_flops = L * 4.0 * batch * H * T * T * 64
    per_kernel = []

    def add(name, entry, work, unit, peak, bound, note):
        ms = by.get(entry)
        if not ms:
            return
        ach = work / ms / (1e9 if unit == "TFLOP/s" else 1e6)
        per_kernel.append({"kernel": name, "ms_per_step": round(ms, 3), "launches_per_step": calls[entry], "bound": bound,
                           "achieved": round(ach, 1), "peak": peak, "unit": unit, "frac": round(ach / peak, 3), "work": note})
    add("attention_fwd_tc_kernel", "b200vit_attention_fwd", attn_fwd_flops, "TFLOP/s", sustained, "tensor",
        "4*B*H*T^2*64 FLOP per layer")
    add("attention_bwd_tc4_kernel", "b200vit_attention_bwd_bias", 2.5 * attn_fwd_flops, "TFLOP/s", sustained, "tensor",
        "2.5x the forward (S, dP, dV, dK, dQ)")
    add("layernorm_fwd_kernel", "b200vit_layernorm_fwd", 2 * L * R * D * 6.0, "GB/s", hbm, "hbm",
        "6 B/element: read f32, write bf16 (2 per layer)")
    add("layernorm_bwd_split_kernel", "b200vit_layernorm_bwd", 2 * L * R * D * 16.0, "GB/s", hbm, "hbm",
        "16 B/element: read bf16 dy, f32 x, f32 residual grad; write f32 grad + bf16 next operand")
    n_params = 85_800_963 if classes == 3 else 86_567_656
    add("adam_kernel", "b200vit_adam_step", n_params * 30.0, "GB/s", hbm, "hbm", "30 B/parameter incl. bf16 shadow refresh")
    ach = gemm_flops / gemm_ms / 1e9
    shape_ms = {}
    for op, epi, M, N, K_, ms in g:
        key = f"op{op}_epi{epi}_{M}x{N}x{K_}"
        a = shape_ms.setdefault(key, [0.0, 0])
        a[0] += ms
        a[1] += 1
    per_shape = {k: {"us": round(v[0] / v[1] * 1e3, 1),
                     "tflops": round(2.0 * float(k.split("_")[2].split("x")[0]) * float(k.split("x")[1]) * float(k.split("x")[2]) / (v[0] / v[1]) / 1e9, 1),
                     "launches_per_step": v[1] / n_steps} for k, v in shape_ms.items()}
    return {
        "bound": "tensor", "achieved": round(ach, 1), "peak": sustained, "unit": "TFLOP/s", "frac": round(ach / sustained, 4),
        "kernel": "gemm_tcgen05_2cta_kernel / gemm_tcgen05_kernel (all %d GEMM launches of a step, FLOP-weighted, timed "
                  "inside real steps)" % round(n_launch),
        "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (kernel timed inside a long step)"
        if "bf16_tflops_sustained" in peaks else "B200_PROFILING.md fallback 1.4 PFLOP/s sustained",
        "flops_per_launch": round(gemm_flops / n_launch), "us_per_launch": round(gemm_ms / n_launch * 1e3, 2),
        "gemm_ms_per_step": round(gemm_ms, 3),
        "all_kernels_ms_per_step": round(sum(by.values()), 3),
        "c_abi_calls_per_step": round(sum(calls.values()), 1),
        "per_shape_in_step": per_shape,
        "per_kernel": per_kernel,
    }


def measure_gemm_roofline(batch, dev, peaks):
    """Replay every GEMM launch shape of one step and time each with CUDA events on the launching stream."""
    total_flops = 0.0
    total_ms = 0.0
    detail = {}
    for name, op, epi, M, N, Kd, count in gemm_shapes_per_step(batch):
        run = make_gemm_runner(name, op, epi, M, N, Kd, dev)
        for _ in range(2):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        fl = 2.0 * M * N * Kd
        detail[name] = {"us": round(ms * 1e3, 1), "tflops": round(fl / ms / 1e9, 1), "launches_per_step": count}
        total_flops += fl * count
        total_ms += ms * count
        del run
    achieved = total_flops / total_ms / 1e9
    peak = peaks.get("bf16_tflops") or 1600.0
    traffic, traffic_src = gemm_traffic_per_launch(batch)
    n_launch = sum(c for *_, c in gemm_shapes_per_step(batch))
    return {
        "achieved": round(achieved, 1), "peak": peak, "unit": "TFLOP/s", "frac": round(achieved / peak, 4),
        "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst: each shape replayed alone, back to back)"
        if "bf16_tflops" in peaks else "B200_PROFILING.md fallback 1.6 PFLOP/s burst",
        "us_per_launch": round(total_ms / n_launch * 1e3, 2), "gemm_ms_per_step": round(total_ms, 3),
        "traffic": None if traffic is None else round(traffic),
        "traffic_unit": "DRAM bytes (read + write) per launch, mean over the step's %d GEMM launches" % n_launch,
        "traffic_source": traffic_src, "per_shape": detail,
    }


# ---------------------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        # nvidia-smi numbers GPUs without regard to CUDA_VISIBLE_DEVICES: address the device by its UUID
        self.index = str(index)
        try:
            self.index = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
        except Exception:
            pass
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        # median over the samples taken under load (upper half of the clock samples)
        busy = sorted(sm)[len(sm) // 2:] if sm else []
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------------
# CPU baseline = the unmodified reference's engine.train_step (oracle/reference_runner.py); the oracle port only if
# the reference copy (baseline/_ref, oracle/install_reference.py) did not travel
# ---------------------------------------------------------------------------------------------------------
def cpu_port_steps(steps: int, warmup: int, batch: int = 32):
    """Fallback: the oracle restatement of engine.train_step (reference recipe, dropout as shipped) on the host cores."""
    from oracle import vit_oracle as O
    from b200vit.models.vit import ViT        # constructor only (parameter init); no b200vit compute runs here
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(42)
    sd = ViT(num_classes=3).state_dict()
    cfg = O.ViTConfig(num_classes=3)
    tr = O.OracleTrainer(sd, cfg, total_steps=1000)
    g = torch.Generator().manual_seed(0)
    X = torch.rand(batch, 3, 224, 224, generator=g)
    y = torch.randint(0, 3, (batch,), generator=g)
    for _ in range(warmup):
        tr.step(X, y, training_dropout=True)
    t0 = time.perf_counter()
    for _ in range(steps):
        tr.step(X, y, training_dropout=True)
    dt = time.perf_counter() - t0
    return {"value": round(batch * steps / dt, 3), "s_per_step": round(dt / steps, 3)}


def cpu_reference_steps(steps: int, warmup: int, batch: int = 32):
    """images/sec of the reference train_step on the host cores: BASELINE.json configs[0] (ViT-B/16, 3 classes, batch 32,
    fp32, dropout as shipped), all host threads."""
    from oracle import reference_runner as RR
    cores = os.cpu_count() or 1
    sample = f"{steps} train steps of batch {batch} (ViT-B/16 3-class fp32, dropout 0.1/0.1/0, after {warmup} warm-up)"
    if RR.reference_root() is not None:
        r = RR.time_reference_train_step("cpu", batch, steps, warmup, threads=cores)
        kind, how = "reference", f"unmodified reference engine.train_step + models/vit.py from {r['reference_root']}"
    else:
        r = cpu_port_steps(steps, warmup, batch)
        kind, how = "port", "oracle/vit_oracle.py OracleTrainer (the reference copy baseline/_ref is missing)"
    return {"value": r["value"], "unit": UNIT, "cores": cores, "kind": kind, "sample": sample, "how": how,
            "s_per_step": r["s_per_step"]}


def gpu_eager_baseline(batch: int, dev_str: str):
    """The unmodified reference model and engine on THIS GPU in eager PyTorch (what a reference user gets from
    `device='cuda'`): fp32 with TF32 off -- the reference's own precision -- and under bf16 autocast."""
    from oracle import reference_runner as RR
    if RR.reference_root() is None:
        return {"unavailable": "reference copy baseline/_ref is missing (oracle/install_reference.py)"}
    out = {"batch": batch, "how": "unmodified reference engine.train_step + models/vit.py, eager PyTorch, wall clock"}
    for key, ac, steps in (("fp32_tf32_off", False, 2), ("bf16_autocast", True, 3)):
        try:
            r = RR.time_reference_train_step(dev_str, batch, steps, 1, autocast=ac)
            out[key] = {"value": r["value"], "unit": UNIT, "ms_per_step": round(r["s_per_step"] * 1e3, 1)}
        except Exception as e:   # e.g. out of memory next to a busy allocator: report, do not kill the bench line
            out[key] = {"error": f"{type(e).__name__}: {str(e)[:160]}"}
        torch.cuda.empty_cache()
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 3))
    warmup = 1 if args.warmup > 0 else 0
    cb = cpu_reference_steps(steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": round(cb["s_per_step"] * 1e3, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ViT-B/16 3-class training step (reference recipe) on host cores, bounded sample: batch 32 "
                               "per step (BASELINE.json configs[0])", "batch_per_step": 32},
        "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "how")},
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def recipe(model, total, FusedAdam):
    decay, no_decay = [], []
    for name, p in model.named_parameters():
        (no_decay if (p.ndim == 1 or name.endswith(".bias")) else decay).append(p)
    opt = FusedAdam([{"params": decay, "weight_decay": 0.03}, {"params": no_decay, "weight_decay": 0.0}], lr=1e-3)
    warm_n = max(1, int(0.05 * total))
    sched = torch.optim.lr_scheduler.SequentialLR(opt, schedulers=[
        torch.optim.lr_scheduler.LinearLR(opt, start_factor=1e-6, end_factor=1.0, total_iters=warm_n),
        torch.optim.lr_scheduler.LinearLR(opt, start_factor=1.0, end_factor=0.0, total_iters=total - warm_n)],
        milestones=[warm_n])
    return opt, sched


def run_extra_config(name, model_kw, image, classes, micro_batch, accum, steps, warmup, world, rank, dev, timers):
    """One of the other BASELINE.json configs as its own small run: images/sec over `steps` optimizer steps of
    `accum` micro-batches of `micro_batch` images per GPU (device-resident synthetic batches, dropout as shipped)."""
    import torch.distributed as dist
    from b200vit.going_modular import engine
    from b200vit.models.vit import ViT
    from b200vit.nn import CrossEntropyLoss
    from b200vit.optim import FusedAdam
    barrier, max_over_ranks = timers
    torch.manual_seed(42)
    model = ViT(num_classes=classes, **model_kw).to(dev)
    net = model
    if world > 1:
        from b200vit.dp import DataParallel
        net = DataParallel(model)
    opt, sched = recipe(model, 10 * (steps + warmup), FusedAdam)
    if world > 1:
        opt.grad_scale = net.grad_scale
    loss_fn = CrossEntropyLoss()
    g = torch.Generator().manual_seed(99 + rank)
    pool = [(torch.rand(micro_batch, 3, image, image, generator=g).to(dev),
             torch.randint(0, classes, (micro_batch,), generator=g).to(dev)) for _ in range(2)]
    engine.train_step(net, [pool[i % 2] for i in range(warmup * accum)], loss_fn, opt, sched, dev, accumulation_steps=accum)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loss, _ = engine.train_step(net, [pool[i % 2] for i in range(steps * accum)], loss_fn, opt, sched, dev,
                                accumulation_steps=accum)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    images = world * micro_batch * accum * steps
    _, gf = train_gflop_per_image(image=image, layers=model_kw.get("num_transformer_layer", 12),
                                  d=model_kw.get("embedding_dim", 768), heads=model_kw.get("num_heads", 12),
                                  mlp=model_kw.get("mlp_size", 3072), classes=classes)
    val = images / (ms / 1e3)
    peak_mem = torch.cuda.max_memory_allocated(dev) / 2 ** 30
    del model, net, opt, sched, pool
    torch.cuda.empty_cache()
    return {"workload": name, "value": round(val, 1), "unit": UNIT, "global_batch": world * micro_batch * accum,
            "per_gpu_micro_batch": micro_batch, "accumulation_steps": accum, "optimizer_steps": steps,
            "ms_per_optimizer_step": round(ms / steps, 2), "train_gflop_per_image": round(gf, 2),
            "mfu_vs_nominal_2250": round(val * gf / 1e3 / world / NOMINAL_BF16_TFLOPS, 4),
            "final_train_loss": round(loss, 4), "peak_mem_gib": round(peak_mem, 1)}


# ---------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200vit", choices=["b200vit", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="per-GPU batch")
    ap.add_argument("--classes", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-eager-baseline", action="store_true")
    ap.add_argument("--no-extra-configs", action="store_true", help="skip the config 3 / config 4 runs")
    ap.add_argument("--config4", action="store_true", help="run ViT-L/16-384 (configs[3]) even when N != 8")
    args = ap.parse_args()

    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch.distributed as dist
    from b200vit import _C
    from b200vit.going_modular import engine
    from b200vit.models.vit import ViT
    from b200vit.nn import CrossEntropyLoss
    from b200vit.optim import FusedAdam

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        sys.exit("bench.py needs a CUDA device: b200vit has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W = max(3, args.warmup)
    K_steps = args.steps
    B = args.batch

    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        try:
            peaks = json.loads(pk.read_text())
        except Exception:
            peaks = {}

    torch.manual_seed(42)
    model = ViT(num_classes=args.classes).to(dev)      # dropout as shipped: 0.1 / 0.1 / 0
    net = model
    if world > 1:
        from b200vit.dp import DataParallel
        net = DataParallel(model)
    decay, no_decay = [], []
    for name, p in model.named_parameters():
        (no_decay if (p.ndim == 1 or name.endswith(".bias")) else decay).append(p)
    opt = FusedAdam([{"params": decay, "weight_decay": 0.03}, {"params": no_decay, "weight_decay": 0.0}], lr=1e-3)
    if world > 1:
        opt.grad_scale = net.grad_scale
    total = 3 * (W + K_steps) + 16
    warm_n = int(0.05 * total)
    sched = torch.optim.lr_scheduler.SequentialLR(opt, schedulers=[
        torch.optim.lr_scheduler.LinearLR(opt, start_factor=1e-6, end_factor=1.0, total_iters=max(1, warm_n)),
        torch.optim.lr_scheduler.LinearLR(opt, start_factor=1.0, end_factor=0.0, total_iters=total - warm_n)],
        milestones=[max(1, warm_n)])
    loss_fn = CrossEntropyLoss()

    g = torch.Generator().manual_seed(1234 + rank)
    n_pool = 4   # distinct host batches, cycled
    host = [(torch.rand(B, 3, 224, 224, generator=g).pin_memory(), torch.randint(0, args.classes, (B,), generator=g).pin_memory())
            for _ in range(n_pool)]
    resident = [(x.to(dev), y.to(dev)) for x, y in host]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    # ---- device-resident run ("value") ---------------------------------------------------------------
    engine.train_step(net, [resident[i % n_pool] for i in range(W)], loss_fn, opt, sched, dev)
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    l0 = _C.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    train_loss, train_acc = engine.train_step(net, [resident[i % n_pool] for i in range(K_steps)], loss_fn, opt, sched, dev)
    e1.record()
    barrier()
    launches = _C.launch_count - l0
    clocks = sampler.stop() if rank == 0 else None
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    ms_step = ms_total / K_steps
    value = world * B * K_steps / (ms_total / 1e3)

    # ---- end-to-end run: host batches, per-step H2D and per-step D2H of the metrics ---------------------
    # one public-API call over K host batches: every step copies its batch H2D from pinned memory (prefetched on a
    # side stream by the engine) and copies its (loss, accuracy) pair D2H to pinned memory (non-blocking, 8 bytes)
    engine.train_step(net, [host[i % n_pool] for i in range(2)], loss_fn, opt, sched, dev)
    barrier()
    e0.record()
    engine.train_step(net, [host[i % n_pool] for i in range(K_steps)], loss_fn, opt, sched, dev)
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    e2e_value = world * B * K_steps / (ms_e2e / 1e3)
    h2d = B * 3 * 224 * 224 * 4 + B * 8
    d2h = 8

    # the same with uint8 host images (the reference's ToDtype(float32, scale=True) moved onto the device): informational
    host_u8 = [((x * 255).to(torch.uint8).pin_memory(), y) for x, y in host]
    engine.train_step(net, [host_u8[i % n_pool] for i in range(2)], loss_fn, opt, sched, dev)
    barrier()
    e0.record()
    engine.train_step(net, [host_u8[i % n_pool] for i in range(K_steps)], loss_fn, opt, sched, dev)
    e1.record()
    barrier()
    ms_u8 = max_over_ranks(e0.elapsed_time(e1))
    del host_u8

    # ---- in-step kernel timing (CUDA events around every C-ABI call of two more real steps; every rank runs them) ----
    step_roof = None
    if not args.no_roofline:
        step_roof = measure_step_kernels(
            lambda n: engine.train_step(net, [resident[i % n_pool] for i in range(n)], loss_fn, opt, sched, dev),
            2, B, peaks, args.classes)
    barrier()
    calls0 = _C.call_count
    engine.train_step(net, [resident[0]], loss_fn, opt, sched, dev)
    c_calls = _C.call_count - calls0
    barrier()

    # ---- where the weak-scaling loss comes from: the same K steps WITHOUT the gradient exchange (no_sync: every rank steps
    # on its own gradients -- a diagnostic, the weights diverge, nothing after this uses the model). The engine's
    # once-per-epoch metric all-reduce at the end lines the ranks up, so every rank measures the SLOWEST rank's time:
    # what N independent GPUs of this box would give at best. ms_per_step minus it is the cost of the exchange.
    dp_diag = None
    if world > 1:
        barrier()
        with net.no_sync():
            engine.train_step(net, [resident[i % n_pool] for i in range(2)], loss_fn, opt, sched, dev)
            barrier()
            e0.record()
            engine.train_step(net, [resident[i % n_pool] for i in range(K_steps)], loss_fn, opt, sched, dev)
            e1.record()
            barrier()
        free_ms = max_over_ranks(e0.elapsed_time(e1)) / K_steps
        dp_diag = {"transport": "peer" if getattr(net, "peer", None) is not None else "nccl",
                   "ms_per_step_without_exchange": round(free_ms, 3), "exchange_cost_ms": round(ms_step - free_ms, 3),
                   "how": "same steps under no_sync (no gradient exchange), slowest rank"}

    # ---- the other BASELINE.json configs (every rank takes part) ------------------------------------------------------
    del host, resident
    extra = {}
    if not args.no_extra_configs:
        del net, opt, sched
        model = None
        torch.cuda.empty_cache()
        timers = (barrier, max_over_ranks)
        accum = max(1, 4096 // (512 * world))
        extra["config3"] = run_extra_config(
            f"BASELINE.json configs[2]: ViT-B/16 1000-class, global batch 4096 = {world} GPU x {accum} micro-batches x 512 "
            f"(gradient accumulation under no_sync, one all-reduce per optimizer step)", {}, 224, 1000, 512, accum,
            steps=3, warmup=1, world=world, rank=rank, dev=dev, timers=timers)
        if world == 8 or args.config4:
            extra["config4"] = run_extra_config(
                f"BASELINE.json configs[3]: ViT-L/16 at 384x384 (577 tokens), 64 images per GPU x {world}",
                dict(image_size=384, num_transformer_layer=24, embedding_dim=1024, num_heads=16, mlp_size=4096),
                384, 1000, 64, 1, steps=5, warmup=2, world=world, rank=rank, dev=dev, timers=timers)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # a run whose loss left the plausible band measured nothing worth reporting (3 random classes: ln 3 = 1.10; the
    # notebook recipe at lr 1e-3 spikes to ~3.5 right after warm-up)
    if not (train_loss == train_loss and 0.2 < train_loss < 10.0):
        sys.exit(f"bench.py: training loss {train_loss} is outside (0.2, 10): the timed run is not a valid training run")

    _, train_gf = train_gflop_per_image(classes=args.classes)
    tflops = value * train_gf / 1e3
    roof = None
    if not args.no_roofline:
        torch.cuda.empty_cache()
        roof = step_roof
        iso = measure_gemm_roofline(B, dev, peaks)
        roof["traffic"] = iso.pop("traffic")
        roof["traffic_unit"] = iso.pop("traffic_unit")
        roof["traffic_source"] = iso.pop("traffic_source")
        roof["gemm_share_of_step"] = round(roof["gemm_ms_per_step"] / ms_step, 3)
        roof["isolated"] = iso
    eager = None
    if not args.no_eager_baseline and world == 1:
        torch.cuda.empty_cache()
        eager = gpu_eager_baseline(B, f"cuda:{local}")
    cb = None
    if not args.no_cpu_baseline and world == 1:
        cb = cpu_reference_steps(steps=2, warmup=1)
        cb.pop("s_per_step", None)

    line = {
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": K_steps, "warmup": W,
        "ms_per_step": round(ms_step, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"ViT-B/16 {args.classes}-class bf16 training, batch {B} per GPU x {world} B200 "
                               f"(BASELINE.json configs[1]{'' if world == 1 else ', weak-scaled data parallel'})",
                   "global_batch": B * world, "tokens": 197, "parallelism": f"dp{world}",
                   "dropout": "0.1/0.1/0 as shipped", "optimizer": "FusedAdam (Adam, coupled L2 0.03/0) + clip 1.0 + per-step LR schedule",
                   "l2": "inputs and activations per step (>1 GB) exceed the 126 MB L2; no explicit flush"},
        "mfu": {"train_gflop_per_image": round(train_gf, 3), "achieved_tflops_per_gpu": round(tflops / world, 1),
                "vs_nominal_2250": round(tflops / world / NOMINAL_BF16_TFLOPS, 4),
                "vs_measured_sustained": round(tflops / world / peaks["bf16_tflops_sustained"], 4) if "bf16_tflops_sustained" in peaks else None},
        "e2e": {"value": round(e2e_value, 1), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(ms_e2e / K_steps, 3)},
        "e2e_uint8_images": {"value": round(world * B * K_steps / (ms_u8 / 1e3), 1), "unit": UNIT,
                             "h2d_bytes_per_step": B * 3 * 224 * 224 + B * 8, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
        "c_abi_calls_per_step": c_calls,
        "clocks": clocks,
        "roofline": roof,
        "cpu_baseline": cb,
        "gpu_eager_baseline": eager,
        "final_train_loss": round(train_loss, 4),
        "loss_check": {"finite": True, "band": [0.2, 10.0]},
    }
    if dp_diag is not None:
        line["data_parallel"] = dp_diag
    line.update(extra)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
