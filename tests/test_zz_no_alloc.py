"""g16_compute_h_noalloc_dev: computeH whose only scratch is the library's static device array (no cudaMalloc).

Collected last on purpose.  The entry point has no counterpart in the reference;
it is checked against the existing g16_compute_h_dev through the C ABI, bit for bit (integer arithmetic: the
tolerance of every parity test in this suite is zero), on calls in a row that share the static array.
"""
import numpy as np
import pytest

from gnark_icicle_b200 import lib

pytestmark = pytest.mark.gpu

FR_TOP_MASK = 0x0FFFFFFFFFFFFFFF   # any 252-bit value is a valid Montgomery Fr image


def _rand_fr(torch, n, gen):
    x = torch.randint(-(2 ** 63), 2 ** 63 - 1, (n, 4), dtype=torch.int64, device="cuda:0", generator=gen)
    x[:, 3] &= FR_TOP_MASK
    return x


def _reference_h(torch, ctx, a, b, c, n):
    """the existing entry point: zero-padded copies (it clobbers its inputs), h left in the first array"""
    pad = [torch.zeros((n, 4), dtype=torch.int64, device="cuda:0") for _ in range(3)]
    for dst, src in zip(pad, (a, b, c)):
        dst[: len(src)] = src
    torch.cuda.synchronize()
    ctx.compute_h_dev(pad[0].data_ptr(), pad[1].data_ptr(), pad[2].data_ptr(), n)
    return pad[0]


def test_noalloc_compute_h_matches_compute_h_dev(ctx):
    import torch
    gen = torch.Generator(device="cuda:0")
    gen.manual_seed(20260501)
    # two calls in a row and then some: a large domain first so that the second, smaller call finds the static array
    # full of stale data beyond its own padding; then the maximum size, an unpadded domain and the tiny ones
    cases = [(1 << 12, 4000), (1 << 10, 700), (1 << 12, 4000), (1 << 16, 60001), (1 << 8, 1 << 8), (4, 3), (2, 1), (1, 1)]
    for n, nc in cases:
        a, b, c = (_rand_fr(torch, nc, gen) for _ in range(3))
        keep = [t.clone() for t in (a, b, c)]
        h = torch.full((n, 4), -1, dtype=torch.int64, device="cuda:0")
        torch.cuda.synchronize()
        ctx.compute_h_noalloc_dev(a.data_ptr(), b.data_ptr(), c.data_ptr(), nc, n, h.data_ptr())
        want = _reference_h(torch, ctx, a, b, c, n)
        torch.cuda.synchronize()
        assert torch.equal(h, want), f"n={n} nc={nc}: no-alloc computeH differs from g16_compute_h_dev"
        for t, k in zip((a, b, c), keep):
            assert torch.equal(t, k), "inputs must be left untouched"


def test_noalloc_rejects_sizes_beyond_the_static_scratch(ctx):
    import torch
    n = 1 << 17
    x = torch.zeros((n, 4), dtype=torch.int64, device="cuda:0")
    with pytest.raises(lib.G16Error, match="exceeds the static scratch"):
        ctx.compute_h_noalloc_dev(x.data_ptr(), x.data_ptr(), x.data_ptr(), n, n, x.data_ptr())
    with pytest.raises(lib.G16Error, match="power of two"):
        ctx.compute_h_noalloc_dev(x.data_ptr(), x.data_ptr(), x.data_ptr(), 5, 6, x.data_ptr())
    # and the context is still usable afterwards
    ctx.compute_h_noalloc_dev(x.data_ptr(), x.data_ptr(), x.data_ptr(), 4, 4, x[8:].data_ptr())


def test_noalloc_allocates_no_device_memory(ctx):
    """free device memory is the same before and after a call on a warmed-up context (the static array is part of
    the module image, loaded with the library)."""
    import torch
    gen = torch.Generator(device="cuda:0")
    gen.manual_seed(7)
    n = 1 << 14
    a, b, c = (_rand_fr(torch, n, gen) for _ in range(3))
    h = torch.empty((n, 4), dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    ctx.compute_h_noalloc_dev(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, n, h.data_ptr())   # warm-up (module load)
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(3):
        ctx.compute_h_noalloc_dev(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, n, h.data_ptr())
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0, (free0, free1)
