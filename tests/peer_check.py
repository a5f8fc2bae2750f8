"""The peer-memory all-reduce kernel (b200vit_allreduce_peer_f32, csrc/peer_reduce.cu) with all `world` ranks played
by ONE GPU: every rank gets its own bucket, flag words and stream, and the `world` kernels run concurrently (each waits
for the others' flags). Checks the sum against torch, bit-identical results in all buckets, ragged sizes, repeated calls
on the same flag words, late ranks. Run as a script (tests/test_peer_reduce_gpu.py does, in a subprocess, so a protocol
bug -- a 20 s bounded spin, then a trap -- cannot take the test session's CUDA context with it)."""
import ctypes
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-vit-paper-replication_b200"))
from b200vit import _C  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    lib = _C.lib()
    for world in (2, 3, 4, 8):
        streams = [torch.cuda.Stream() for _ in range(world)]
        flags = [torch.zeros(_C.PEER_FLAG_WORDS, dtype=torch.int32, device=dev) for _ in range(world)]
        flag_tab = (ctypes.c_void_p * world)(*[f.data_ptr() for f in flags])
        seq = 0
        for n in (4, 64, 1000, 4 * 12345, 7_000_000):
            g = torch.Generator(device=dev).manual_seed(world * 1000 + n % 997)
            for rep in range(3):
                bufs = [torch.randn(n, device=dev, generator=g) for _ in range(world)]
                want = bufs[0].clone()
                for b in bufs[1:]:
                    want += b                                   # rank order, like the kernel
                buf_tab = (ctypes.c_void_p * world)(*[b.data_ptr() for b in bufs])
                seq += 1
                torch.cuda.synchronize()
                for r in range(world):
                    if (r + rep) % world == 0:
                        with torch.cuda.stream(streams[r]):
                            torch.cuda._sleep(2_000_000)        # this rank arrives late
                    ctas = (8, 148, 296)[rep]
                    _C.check(lib.b200vit_allreduce_peer_f32(buf_tab, flag_tab, world, r, n, seq, ctas,
                                                            streams[r].cuda_stream), "b200vit_allreduce_peer_f32")
                torch.cuda.synchronize()
                for r, b in enumerate(bufs):
                    assert torch.equal(b, want), f"world {world} n {n} rep {rep}: rank {r}'s bucket differs from the sum"
        for f in flags:     # the launch-local counters are back to zero; the signalling words hold the last seq
            assert int(f[32:].abs().sum()) == 0
    # argument errors
    one = torch.zeros(8, device=dev)
    tab = (ctypes.c_void_p * 2)(one.data_ptr(), one.data_ptr())
    assert lib.b200vit_allreduce_peer_f32(tab, tab, 9, 0, 8, 1, 8, None) < 0 and b"world" in lib.b200vit_last_error()
    assert lib.b200vit_allreduce_peer_f32(tab, tab, 2, 0, 6, 1, 8, None) < 0 and b"multiple of 4" in lib.b200vit_last_error()
    assert lib.b200vit_allreduce_peer_f32(tab, tab, 2, 2, 8, 1, 8, None) < 0 and b"rank" in lib.b200vit_last_error()
    print("peer_check ok")


if __name__ == "__main__":
    main()
