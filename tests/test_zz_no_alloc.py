"""b200vit_l2norm_sq_noalloc: the gradient-norm entry point for hosts that cannot allocate device memory
(an arena-owning server where cudaMalloc fails; a loop that cannot take the device-wide stall of cudaMalloc /
cudaFree). It keeps its partial sums in a fixed array inside the library instead of a caller buffer.

Checked here, through the C ABI: on two calls in a row (the second on a different, smaller problem, so anything
stale in the shared built-in scratch would show) it returns exactly what b200vit_l2norm_sq returns with a
caller-provided scratch buffer, and problems beyond its stated limit are refused with a clear message.
This file sorts last on purpose: a `-x` run has executed every other test before it gets here."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

CHUNK = 8192              # elements per partial sum (csrc/optim.cu)
STATIC_PARTIALS = 262144  # capacity of the built-in scratch (include/b200vit.h)


def _tables(tensors):
    dev = tensors[0].device
    ptrs = torch.tensor([t.data_ptr() for t in tensors], dtype=torch.int64, device=dev)
    sizes = torch.tensor([t.numel() for t in tensors], dtype=torch.int64, device=dev)
    return ptrs, sizes, max(t.numel() for t in tensors)


def _both(tensors):
    from b200vit import kernels as K
    dev = tensors[0].device
    ptrs, sizes, mx = _tables(tensors)
    n = len(tensors)
    partials = torch.empty(n * math.ceil(mx / CHUNK), dtype=torch.float32, device=dev)
    ref = torch.zeros(1, device=dev)
    K.l2norm_sq(ptrs, sizes, n, mx, partials, ref)
    got = torch.full((1,), float("nan"), device=dev)
    K.l2norm_sq_noalloc(ptrs, sizes, n, mx, got)
    torch.cuda.synchronize()
    return got.item(), ref.item()


def test_noalloc_norm_matches_scratch_variant_twice():
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(11)
    # call 1: a ViT-B-like mix of big and small tensors, odd sizes included
    big = [torch.randn(s, generator=g).to(dev) for s in [(768, 3072), (2304, 768), (768,), (1, 197, 768), (1001,), (3,)]]
    got1, ref1 = _both(big)
    exact = sum(float((t.double() ** 2).sum()) for t in big)
    assert got1 == ref1                                   # same kernels, same summation order: bit-identical
    assert abs(got1 - exact) <= 1e-5 * exact              # and right (tolerance of the other optimizer tests)
    # call 2, straight after, on a smaller problem: must not see leftovers of call 1 in the shared scratch
    small = [torch.randn(s, generator=g).to(dev) * 3.0 for s in [(100,), (8192,), (8193,), (5, 7)]]
    got2, ref2 = _both(small)
    exact2 = sum(float((t.double() ** 2).sum()) for t in small)
    assert got2 == ref2
    assert abs(got2 - exact2) <= 1e-5 * exact2


def test_noalloc_norm_refuses_more_than_its_scratch_holds():
    from b200vit import kernels as K
    dev = torch.device("cuda:0")
    t = torch.zeros(16, device=dev)
    ptrs, sizes, _ = _tables([t, t])
    out = torch.zeros(1, device=dev)
    too_big = (STATIC_PARTIALS // 2 + 1) * CHUNK           # 2 tensors x that many chunks > capacity
    with pytest.raises(RuntimeError, match="built-in scratch holds 262144"):
        K.l2norm_sq_noalloc(ptrs, sizes, 2, too_big, out)
