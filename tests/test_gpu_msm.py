"""GPU parity: G1/G2 Pippenger MSM and the fixed-base batch multiplication vs the CPU oracle (bit-exact affine)."""
import numpy as np
import pytest

from helpers import (fr_rand, g1_bases, g2_bases, trapdoor_g1, trapdoor_g2, assert_limbs_equal, fr_from_ints,
                     g1_gen_arr, g2_gen_arr, R)
from oracle import cpu as ocpu
from oracle import bn254 as C
from oracle import msm as omsm

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 2, 3, 17, 255, 1000, 4096, 50_000])
def test_msm_g1_matches_oracle(ctx, n):
    k, pts = g1_bases(100 + n, n)
    s = fr_rand(200 + n, n)
    got = ctx.msm_g1(s, pts)
    assert_limbs_equal(got, ocpu.msm_g1(pts, s), f"msm_g1 n={n} vs CPU Pippenger")
    assert_limbs_equal(got, trapdoor_g1(s, k), f"msm_g1 n={n} vs trapdoor identity")


@pytest.mark.parametrize("n", [1, 5, 300, 5000])
def test_msm_g2_matches_oracle(ctx, n):
    k, pts = g2_bases(300 + n, n)
    s = fr_rand(400 + n, n)
    got = ctx.msm_g2(s, pts)
    assert_limbs_equal(got, ocpu.msm_g2(pts, s), f"msm_g2 n={n} vs CPU Pippenger")
    assert_limbs_equal(got, trapdoor_g2(s, k), f"msm_g2 n={n} vs trapdoor identity")


def test_msm_small_against_python_naive(ctx):
    """independent of the C oracle: double-and-add on Python ints."""
    n = 9
    k, pts = g1_bases(77, n)
    s = fr_rand(78, n)
    want = omsm.msm_naive([ocpu.g1_pt(p) for p in pts], ocpu.fr_ints(s))
    assert ocpu.g1_pt(ctx.msm_g1(s, pts)) == want


def test_msm_zero_length_is_infinity(ctx):
    assert not ctx.msm_g1(np.zeros((0, 4), np.uint64), np.zeros((0, 8), np.uint64)).any()
    assert not ctx.msm_g2(np.zeros((0, 4), np.uint64), np.zeros((0, 16), np.uint64)).any()


@pytest.mark.parametrize("c", list(range(8, 17)))
def test_msm_g1_every_window_size(ctx, c):
    n = 3000
    k, pts = g1_bases(500, n)
    s = fr_rand(501, n)
    ctx.set("msm_c", c)
    try:
        assert_limbs_equal(ctx.msm_g1(s, pts), trapdoor_g1(s, k), f"msm_g1 c={c}")
    finally:
        ctx.set("msm_c", 0)


def test_msm_g1_adversarial_scalars(ctx):
    """scalars in {0, 1, r-1, 2^k boundary values}: zero digits, carries through every window, negative digits."""
    n = 4096
    k, pts = g1_bases(600, n)
    special = [0, 1, R - 1, R - 2, 2, (1 << 16) - 1, 1 << 16, (1 << 15), (1 << 15) - 1, (1 << 253), (1 << 128) - 1,
               0x8000800080008000800080008000800080008000800080008000800080008000 % R,
               0x7fff7fff7fff7fff7fff7fff7fff7fff7fff7fff7fff7fff7fff7fff7fff7fff % R,
               0xffffffffffffffffffffffffffffffffffffffffffffffffffffffffffffffff % R]
    vals = [special[i % len(special)] for i in range(n)]
    s = fr_from_ints(vals)
    assert_limbs_equal(ctx.msm_g1(s, pts), trapdoor_g1(s, k), "adversarial scalars")
    # all scalars zero -> infinity
    assert not ctx.msm_g1(np.zeros((n, 4), np.uint64), pts).any()


def test_msm_g1_degenerate_bases(ctx):
    """DummySetup-like keys (setup.go:544-558: every base identical), infinity bases and P/-P pairs."""
    n = 2048
    g = g1_gen_arr()
    s = fr_rand(700, n)
    # (a) all bases equal: every bucket add is a doubling
    k1 = fr_rand(701, 1)
    p = ocpu.g1_mul(g, k1[0])
    pts = np.repeat(p.reshape(1, 8), n, axis=0)
    assert_limbs_equal(ctx.msm_g1(s, pts), trapdoor_g1(s, np.repeat(k1, n, axis=0)), "all-equal bases")
    # (b) all scalars equal too: one bucket per window gets everything (long bucket split across threads)
    s1 = np.repeat(fr_rand(702, 1), n, axis=0)
    assert_limbs_equal(ctx.msm_g1(s1, pts), trapdoor_g1(s1, np.repeat(k1, n, axis=0)), "one heavy bucket")
    # (c) infinity bases are skipped; P and -P with the same scalar cancel
    k, pts2 = g1_bases(703, n)
    pts2 = pts2.copy()
    pts2[::7] = 0
    kk = ocpu.fr_ints(k)
    kk = [0 if i % 7 == 0 else v for i, v in enumerate(kk)]
    neg = ocpu.g1_arr([C.g1_neg(ocpu.g1_pt(pts2[1]))])[0]
    pts2[2] = neg
    kk[2] = (-kk[1]) % R
    s2 = s.copy()
    s2[2] = s2[1]
    assert_limbs_equal(ctx.msm_g1(s2, pts2), trapdoor_g1(s2, fr_from_ints(kk)), "infinity + cancelling bases")


def test_msm_g1_skewed_scalars(ctx):
    """witness-like distribution: mostly 0/1/small values (SURVEY.md hard part 1)."""
    n = 60_000
    k, pts = g1_bases(800, n)
    rng = np.random.default_rng(5)
    vals = [int(v) for v in rng.choice([0, 1, 1, 1, 2, 3, 255], size=n)]
    big = ocpu.fr_ints(fr_rand(801, n // 50))
    for i, v in enumerate(big):
        vals[i * 50] = v
    s = fr_from_ints(vals)
    assert_limbs_equal(ctx.msm_g1(s, pts), trapdoor_g1(s, k), "skewed scalars")


def test_msm_heavy_buckets_are_reduced_cooperatively(ctx):
    """one bucket per window holds everything (all scalars equal): thousands of slice pieces per bucket go through
    k_fixup_heavy; also every window size must cope with the 2-bit top window of real Fr scalars."""
    n = 150_000
    k = fr_rand(950, n)
    pts = ctx.g1_fixed_base_mul(k)
    s1 = np.repeat(fr_rand(951, 1), n, axis=0)
    assert_limbs_equal(ctx.msm_g1(s1, pts), trapdoor_g1(s1, k), "all-equal scalars, distinct bases")
    s = fr_rand(952, n)
    for c in (9, 12, 14):
        ctx.set("msm_c", c)
        try:
            assert_limbs_equal(ctx.msm_g1(s, pts), trapdoor_g1(s, k), f"c={c} short top window")
        finally:
            ctx.set("msm_c", 0)
    m = 20_000
    p2 = ctx.g2_fixed_base_mul(k[:m])
    assert_limbs_equal(ctx.msm_g2(s1[:m], p2), trapdoor_g2(s1[:m], k[:m]), "G2 all-equal scalars")


def test_msm_g2_degenerate(ctx):
    n = 512
    g = g2_gen_arr()
    k1 = fr_rand(900, 1)
    p = ocpu.g2_mul(g, k1[0])
    pts = np.repeat(p.reshape(1, 16), n, axis=0)
    s = fr_rand(901, n)
    s[::3] = 0
    assert_limbs_equal(ctx.msm_g2(s, pts), trapdoor_g2(s, np.repeat(k1, n, axis=0)), "G2 all-equal bases")


@pytest.mark.parametrize("n", [1, 100, 20_000])
def test_fixed_base_mul_matches_oracle(ctx, n):
    k = fr_rand(1000 + n, n)
    k[0] = 0
    assert_limbs_equal(ctx.g1_fixed_base_mul(k), ocpu.g1_mul_gen_batch(k), "g1 fixed base")
    m = min(n, 3000)
    assert_limbs_equal(ctx.g2_fixed_base_mul(k[:m]), ocpu.g2_mul_gen_batch(k[:m]), "g2 fixed base")


def test_msm_g1_large_trapdoor(ctx):
    """BASELINE config 2 scale (2^22 here to keep the suite short): bases from the GPU fixed-base kernel
    (itself checked against the oracle above), result via the trapdoor identity; bit-exact."""
    n = 1 << 22
    k = fr_rand(1100, n)
    pts = ctx.g1_fixed_base_mul(k)
    s = fr_rand(1101, n)
    assert_limbs_equal(ctx.msm_g1(s, pts), trapdoor_g1(s, k), "msm_g1 2^22")


def test_msm_g2_2p20_trapdoor(ctx):
    """BASELINE config 2: G2 MSM at 2^20."""
    n = 1 << 20
    k = fr_rand(1200, n)
    pts = ctx.g2_fixed_base_mul(k)
    s = fr_rand(1201, n)
    assert_limbs_equal(ctx.msm_g2(s, pts), trapdoor_g2(s, k), "msm_g2 2^20")


# ---------------------------------------------------------------- resident bases with precomputed multiples
@pytest.mark.parametrize("n,g2", [(1 << 16, False), (50000, False), (1 << 14, True), (1000, False), (0, False)])
def test_msm_resident_bases(ctx, n, g2):
    """g16_bases_upload keeps the bases (and, from 2^14 points, the table of 2^(c j) multiples) on the device;
    g16_msm_resident_dev must give the same point as the plain MSM and as the trapdoor identity."""
    import torch
    k = fr_rand(77 + n % 13, n)
    s = fr_rand(78 + n % 13, n)
    if n:
        s[:3] = fr_from_ints([0, 1, R - 1])
    bases = (ctx.g2_fixed_base_mul(k) if g2 else ctx.g1_fixed_base_mul(k)) if n else np.zeros((0, 16 if g2 else 8), dtype=np.uint64)
    if n > 10:
        bases[7] = 0                       # an infinity base is legal
        k = k.copy(); k[7] = 0
    h = ctx.bases_upload(bases, g2=g2)
    try:
        sd = torch.from_numpy(np.ascontiguousarray(s).view(np.int64)).cuda() if n else torch.zeros((1, 4), dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        got = ctx.msm_resident_dev(h, sd.data_ptr(), g2=g2)
        plain = (ctx.msm_g2 if g2 else ctx.msm_g1)(s, bases)
        assert_limbs_equal(got, plain, "resident vs plain MSM")
        if n:
            assert_limbs_equal(got, (trapdoor_g2 if g2 else trapdoor_g1)(s, k), "resident MSM vs trapdoor")
        else:
            assert not got.any()
    finally:
        ctx.bases_free(h)
