"""A multi-GPU context behind the single-call C ABI (SURVEY.md 8(b) + 8(e)): g16_ctx_create over several devices,
g16_pk_upload placing the key itself, and ONE g16_prove_bn254 / _dev / _wires call returning the combined five points.

On a one-GPU box the group's ranks are emulated on device 0 (the same ordinal listed several times): the whole
choreography runs -- rank threads, key placement on the cost line, the D(rb) loads, the three exchange kernels with
their cross-rank events, the pulls of h under every G1.Z slice, the XYZZ sum of the partial results -- only the peer
mapping is trivial.  With >= 2 GPUs visible the same tests also run over real devices.  Everything is compared bit for
bit with the single-GPU context AND with the oracle's trapdoor expectation."""
import numpy as np
import pytest

from gnark_icicle_b200 import lib
import gnark_icicle_b200 as g16
from oracle import cpu as ocpu

from helpers import ChainCase, chain_csr, assert_limbs_equal

pytestmark = pytest.mark.gpu


def _devices(world):
    import torch
    have = torch.cuda.device_count()
    return list(range(world)) if have >= world else [0] * world


def _upload(c, case):
    pk = case.pk
    return c.pk_upload(pk.domain_n, pk.g1_a, pk.g1_b, pk.g1_k, pk.g1_z, pk.g2_b, pk.infinity_a, pk.infinity_b, pk.nb_wires)


@pytest.fixture(scope="module")
def case():
    c = ChainCase(20000, seed=61)          # n = 2^15: every rank of an 8-GPU group keeps 2^12 elements
    c.h = ocpu.compute_h(c.a, c.b, c.c, c.n)
    c.want = c.expected_points(c.h)
    return c


@pytest.mark.parametrize("world", [2, 4, 8])
def test_group_prove_equals_single_gpu_and_oracle(ctx, case, world):
    import torch
    single = _upload(ctx, case)
    ref, _ = ctx.prove(single, case.W, case.a, case.b, case.c)
    ctx.pk_free(single)
    assert_limbs_equal(ref, case.want, "single-GPU prove vs trapdoor")
    grp = lib.Context(_devices(world))
    try:
        h = _upload(grp, case)
        # the placement tiles every base array exactly once
        lens = [len(case.pk.g1_a), len(case.pk.g1_b), len(case.pk.g1_k), case.n - 1, len(case.pk.g2_b)]
        for i in range(5):
            pos = 0
            for r in range(world):
                lo, hi = grp.pk_placement(h, r)[i]
                assert lo == hi or lo == pos, (i, r, lo, hi, pos)
                pos = hi if hi > lo else pos
            assert pos == lens[i], (i, pos, lens[i])
        # host buffers (twice: the second call finds every workspace and exchange buffer warm and dirty)
        for _ in range(2):
            pts, st = grp.prove(h, case.W, case.a, case.b, case.c)
            assert_limbs_equal(pts, case.want, f"group of {world}: host-buffer prove")
        assert st["kernel_launches"] > 0 and sum(st["msm_n"]) == sum(lens)
        # device-resident inputs on the group's first GPU
        dev = torch.device("cuda:0")
        Wd = torch.from_numpy(case.W.view(np.int64)).to(dev)
        abc = torch.zeros((3, case.n, 4), dtype=torch.int64, device=dev)
        for i, x in enumerate((case.a, case.b, case.c)):
            abc[i, :len(x)] = torch.from_numpy(x.view(np.int64)).to(dev)
        torch.cuda.synchronize()
        pts, _ = grp.prove_dev(h, Wd.data_ptr(), abc[0].data_ptr(), abc[1].data_ptr(), abc[2].data_ptr())
        assert_limbs_equal(pts, case.want, f"group of {world}: device-resident prove")
        # wire vector only: every rank evaluates its rows of the resident constraint matrices
        cs = chain_csr(case.nc, _cseed(61))
        r1 = grp.r1cs_upload(cs.nb_constraints, cs.nb_wires, cs.coefficients, (cs.L, cs.R, cs.O))
        pts, stw = grp.prove_wires(h, r1, case.W)
        assert_limbs_equal(pts, case.want, f"group of {world}: wires-only prove")
        assert stw["h2d_bytes"] == case.nw * 32          # every rank uploads one n-th of W, the rest moves over NVLink
        grp.r1cs_free(r1)
        grp.pk_free(h)
    finally:
        grp.close()


def _cseed(seed):
    from oracle.prng import SEED_BASE
    return SEED_BASE + 2000 + seed


def test_group_with_a_domain_too_small_to_distribute(ctx):
    """a tiny circuit on a group context lives on the first GPU; same call, same result"""
    small = ChainCase(100, seed=62)
    want = small.expected_points(ocpu.compute_h(small.a, small.b, small.c, small.n))
    grp = lib.Context(_devices(4))
    try:
        h = _upload(grp, small)
        assert grp.pk_placement(h, 1) == [(0, 0)] * 5
        pts, _ = grp.prove(h, small.W, small.a, small.b, small.c)
        assert_limbs_equal(pts, want, "tiny circuit on a group context")
        grp.pk_free(h)
    finally:
        grp.close()


def test_group_rejects_caller_side_shards_and_bad_device_lists(ctx, case):
    grp = lib.Context(_devices(2))
    try:
        pk = case.pk
        with pytest.raises(lib.G16Error, match="shards the key itself"):
            grp.pk_upload(pk.domain_n, pk.g1_a, pk.g1_b, pk.g1_k, pk.g1_z, pk.g2_b, pk.infinity_a, pk.infinity_b, pk.nb_wires,
                          shard_index=0, shard_count=2)
        with pytest.raises(lib.G16Error, match="unknown proving-key handle"):
            grp.prove(12345, case.W, case.a, case.b, case.c)
    finally:
        grp.close()
    with pytest.raises(lib.G16Error):
        lib.Context([0, 0, 0])           # 3 devices: not a power of two
    with pytest.raises(lib.G16Error):
        lib.Context([0, 99])             # no such GPU


def test_sharded_key_with_wires_only_entry_point(ctx, case):
    """the constraint matrices gather arbitrary wires, so the wires-only entry point must upload the whole
    wire vector even when this context's shard of the key reads a sub-range of it."""
    cs = chain_csr(case.nc, _cseed(61))
    pk = case.pk
    blocks = []
    r1 = ctx.r1cs_upload(cs.nb_constraints, cs.nb_wires, cs.coefficients, (cs.L, cs.R, cs.O))
    for i in range(3):
        # poison the staging buffers with another witness first
        h = ctx.pk_upload(pk.domain_n, pk.g1_a, pk.g1_b, pk.g1_k, pk.g1_z, pk.g2_b, pk.infinity_a, pk.infinity_b, pk.nb_wires,
                          shard_index=i, shard_count=3)
        junk = ocpu.fr_random(900 + i, case.nw)
        ctx.prove_wires(h, r1, junk)
        pts, _ = ctx.prove_wires(h, r1, case.W)
        blocks.append(pts)
        ctx.pk_free(h)
    ctx.r1cs_free(r1)
    assert_limbs_equal(g16.combine_shards(blocks), case.want, "3-way sharded wires-only prove")


def test_prove_surface_over_a_group_and_shard_rejection(case):
    """gnark_icicle_b200.Prove (the mirror of icicle_bn254.Prove) over a multi-GPU context gives the proof the
    single-GPU context gives; a single shard of the key is refused instead of yielding a proof that cannot verify."""
    sol = g16.R1CSSolution(W=case.W, A=case.a, B=case.b, C=case.c)
    opts = (g16.WithIcicleAcceleration(), g16.WithRandomness(case.r, case.s))
    one = g16.Prove(None, case.pk, sol, *opts)
    case.pk.free_device()
    grp = g16.Prove(None, case.pk, sol, g16.WithDevices(_devices(4)), *opts)
    case.pk.free_device()
    assert grp.raw() == one.raw()
    assert grp.raw() == g16.assemble(case.pk, case.want, case.r, case.s).raw()
    with pytest.raises(lib.G16Error, match="needs the whole key"):
        g16.Prove(None, case.pk, sol, g16.WithDevice(0, 0, 2), *opts)


# ---------------------------------------------------------------- group primitives (BASELINE configs 2 and 3 on N GPUs)
FR_TOP_MASK = 0x0FFFFFFFFFFFFFFF


def _rand_fr(torch, n, gen):
    x = torch.randint(-(2 ** 63), 2 ** 63 - 1, (n, 4), dtype=torch.int64, device="cuda:0", generator=gen)
    x[:, 3] &= FR_TOP_MASK
    return x


@pytest.mark.parametrize("world", [2, 8])
def test_group_transforms_and_compute_h_equal_single_gpu(ctx, world):
    """g16_ntt_dev / g16_compute_h_dev on a group context: one exchange per transform between the D(rb) and the block
    distribution, arrays on the group's first GPU -- bit-identical to the single-GPU context (itself checked against
    oracle/c up to 2^26 in test_gpu_large.py)."""
    import torch
    logn = 20
    n = 1 << logn
    gen = torch.Generator(device="cuda:0")
    gen.manual_seed(7100 + world)
    grp = lib.Context(_devices(world))
    try:
        x = _rand_fr(torch, n, gen)
        for kw in (dict(inverse=True, coset=False, decimation=lib.DIF), dict(inverse=True, coset=True, decimation=lib.DIF),
                   dict(inverse=False, coset=True, decimation=lib.DIT), dict(inverse=False, coset=False, decimation=lib.DIT),
                   dict(inverse=False, coset=True, decimation=lib.DIF)):      # the last shape is not distributed: first GPU alone
            a, b = x.clone(), x.clone()
            torch.cuda.synchronize()
            ctx.ntt_dev(a.data_ptr(), n, **kw)
            grp.ntt_dev(b.data_ptr(), n, **kw)
            torch.cuda.synchronize()
            assert torch.equal(a, b), kw
        abc = torch.stack([_rand_fr(torch, n, gen) for _ in range(3)])
        one, many = abc.clone(), abc.clone()
        torch.cuda.synchronize()
        ctx.compute_h_dev(one[0].data_ptr(), one[1].data_ptr(), one[2].data_ptr(), n)
        grp.compute_h_dev(many[0].data_ptr(), many[1].data_ptr(), many[2].data_ptr(), n)
        torch.cuda.synchronize()
        assert torch.equal(one[0], many[0]), "distributed computeH differs from the single-GPU h"
    finally:
        grp.close()


@pytest.mark.parametrize("world", [2, 4])
def test_group_resident_msm_equals_trapdoor(ctx, world):
    """g16_bases_upload on a group context shards the points by range (every slice with its own table of multiples);
    g16_msm_resident(_dev) adds the ranks' partial sums.  Expected value from the exponent side."""
    import torch
    from helpers import g1_gen_arr
    n = 1 << 19
    gen = torch.Generator(device="cuda:0")
    gen.manual_seed(7200 + world)
    k, s = _rand_fr(torch, n, gen), _rand_fr(torch, n, gen)
    bases = torch.empty((n, 8), dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    ctx.g1_fixed_base_mul_dev(k.data_ptr(), n, bases.data_ptr())
    kh, sh = k.cpu().numpy().view(np.uint64), s.cpu().numpy().view(np.uint64)
    want = ocpu.g1_mul(g1_gen_arr(), ocpu.fr_dot(sh, kh))
    grp = lib.Context(_devices(world))
    try:
        h = grp.bases_upload_dev(bases.data_ptr(), n)
        assert_limbs_equal(grp.msm_resident_dev(h, s.data_ptr()), want, "group resident MSM, device scalars")
        assert_limbs_equal(grp.msm_resident(h, sh), want, "group resident MSM, host scalars")
        grp.bases_free(h)
        h = grp.bases_upload(bases.cpu().numpy().view(np.uint64))
        assert_limbs_equal(grp.msm_resident_dev(h, s.data_ptr()), want, "group resident MSM, bases from the host")
        grp.bases_free(h)
    finally:
        grp.close()
