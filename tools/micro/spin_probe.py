"""GEMM time with a small co-resident kernel of another stream (tools/micro/spin.cu) in its different modes."""
import ctypes, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200")); sys.path.insert(0, str(ROOT))
import torch, bench
spin = ctypes.CDLL(str(ROOT / "tools/micro/libspin.so"))
spin.spin_launch.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p]
dev = torch.device("cuda:0")
flag = torch.zeros(64, dtype=torch.int32, device=dev)
buf = torch.zeros(7_000_000, device=dev)
side = torch.cuda.Stream()
shapes = {s[0]: s for s in bench.gemm_shapes_per_step(256)}
MODES = {0: "sleep only", 1: "poll ld.acquire.sys", 2: "poll ld.relaxed.sys", 3: "poll ld.acquire.gpu", 4: "stream ld/st 28 MB"}
for n in ("fc2_wgrad", "fc2_dgrad", "qkv_dgrad"):
    name, op, epi, M, N, Kd, _ = shapes[n]
    run = bench.make_gemm_runner(name, op, epi, M, N, Kd, dev)
    def timed(reps=20):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): run()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps * 1e3
    for _ in range(3): run()
    torch.cuda.synchronize()
    print(f"{n}: alone {timed():.1f} us")
    for mode in (0, 1, 2, 3, 4):
        for ctas, sleep in ((148, 200), (148, 2000), (296, 200)):
            torch.cuda.synchronize()
            rc = spin.spin_launch(mode, ctas, 12_000_000, flag.data_ptr(), buf.data_ptr(), buf.numel() // 4, sleep, side.cuda_stream)
            assert rc == 0
            torch.cuda._sleep(200_000)          # let the spinner settle on the SMs first
            t = timed()
            torch.cuda.synchronize()
            print(f"   with {ctas} CTAs {MODES[mode]:22s} sleep {sleep:4d} ns: {t:.1f} us")
