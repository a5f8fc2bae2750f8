"""perf_timeline.py under torchrun (data parallel, 256 images per GPU): rank 0 prints kernel busy time, wall time, the
NCCL kernels and the largest gaps of 3 steps (not a pytest file)."""
import os, sys, collections
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
import torch, torch.distributed as dist
from torch.profiler import ProfilerActivity, profile
from b200vit.dp import DataParallel
from b200vit.models.vit import ViT
from b200vit.nn import CrossEntropyLoss
from b200vit.optim import FusedAdam
local = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
torch.manual_seed(0)
m = ViT(num_classes=3).to(dev).train()
net = DataParallel(m)
opt = FusedAdam(m.parameters(), lr=1e-4); opt.grad_scale = net.grad_scale
lf = CrossEntropyLoss()
X = torch.rand(256, 3, 224, 224, device=dev); y = torch.randint(0, 3, (256,), device=dev)
def step():
    loss = lf(net(X), y); opt.zero_grad(); loss.backward(); opt.step(max_norm=1.0)
for _ in range(4): step()
torch.cuda.synchronize(); dist.barrier()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): step()
    torch.cuda.synchronize()
import json as _json, tempfile as _tempfile
_tf = _tempfile.mktemp(suffix=".json"); prof.export_chrome_trace(_tf)
_tr = [e for e in _json.load(open(_tf))["traceEvents"] if e.get("cat") == "kernel"]
_tr.sort(key=lambda e: e["ts"])
if os.path.isdir("gpurun_out"):
    with open(f"gpurun_out/comm_rank{dist.get_rank()}.txt", "w") as f:
        for e in _tr:
            f.write(f"{e['ts']:.1f} {e['dur']:.1f} {e['name'].replace('void ', '').replace('b200vit::', '')[:48]}\n")
if dist.get_rank() == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ks = sorted([(e.time_range.start, e.time_range.end, e.name) for e in evs])
    wall = ks[-1][1] - ks[0][0]
    nccl = [(s, e, n) for s, e, n in ks if "nccl" in n.lower() or "allreduce_peer" in n]
    comp = [(s, e, n) for s, e, n in ks if "nccl" not in n.lower() and "allreduce_peer" not in n]
    print(f"wall {wall/3e3:.2f} ms/step; compute-kernel busy {sum(e-s for s,e,_ in comp)/3e3:.2f} ms/step; "
          f"nccl kernels {len(nccl)/3:.0f}/step, {sum(e-s for s,e,_ in nccl)/3e3:.2f} ms/step")
    for s, e, n in nccl[:16]: print(f"   nccl {e-s:8.1f} us  {n[:60]}")
    # idle gaps of the compute stream
    gaps = sorted([(comp[i+1][0] - comp[i][1], comp[i][2][:45], comp[i+1][2][:45]) for i in range(len(comp)-1)], reverse=True)
    print("compute-stream gaps > 20 us: %d, total %.2f ms over 3 steps" % (sum(g[0] > 20 for g in gaps), sum(g[0] for g in gaps if g[0] > 20)/1e3))
    for g in gaps[:10]: print(f"{g[0]:8.1f} us  after {g[1]}  before {g[2]}")
    # compute kernels overlapping an NCCL kernel: duration against the median of the same kernel name + rounded size class
    import statistics
    byname = collections.defaultdict(list)
    for s_, e_, n in comp: byname[n].append(e_ - s_)
    slow = 0.0
    for ns, ne, _ in nccl:
        for s_, e_, n in comp:
            if s_ < ne and e_ > ns and "gemm" in n:
                durs = sorted(byname[n])
                # nearest-duration cluster: the closest 25% quantile below is the unperturbed time of this shape class
                lower = [d for d in durs if d <= (e_ - s_)]
                print(f"   overlap: nccl [{ns-ks[0][0]:9.1f},{ne-ks[0][0]:9.1f}]  {n[28:75]} start {s_-ks[0][0]:9.1f} dur {e_-s_:7.1f} us")
    # everything within 150 us before / 450 us after the start of the 5th NCCL kernel, with launch geometry
    import json, tempfile
    tf = tempfile.mktemp(suffix=".json"); prof.export_chrome_trace(tf)
    tr = [e for e in json.load(open(tf))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
    tr.sort(key=lambda e: e["ts"])
    nk = [e for e in tr if "nccl" in e["name"].lower() or "allreduce_peer" in e["name"]]
    if len(nk) > 5:
        t0 = nk[5]["ts"]
        for e in tr:
            if t0 - 150 <= e["ts"] <= t0 + 450:
                a = e.get("args", {})
                print(f"   win {e['ts']-t0:8.1f} +{e['dur']:7.1f} us  stream {a.get('stream')}  grid {a.get('grid')} block {a.get('block')} regs {a.get('registers per thread')} smem {a.get('shared memory')}  {e['name'][:70]}")
    agg = collections.defaultdict(float)
    for s, e, n in comp: agg[n[:50]] += e - s
    for n, t in sorted(agg.items(), key=lambda kv: -kv[1])[:12]: print(f"{t/3e3:8.3f} ms/step  {n}")
dist.destroy_process_group()
