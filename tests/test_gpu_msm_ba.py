"""GPU: the batch-affine pre-reduction of the bucket entries (csrc/msm_ba.cuh, tunable "msm_ba_rounds") gives exactly the
points the direct XYZZ accumulation gives -- on random inputs and on the inputs that exercise every branch of the
affine pair addition: equal points (doubling; DummySetup-like keys), P and -P in one bucket (cancellation to infinity),
bases at infinity, zero scalars, 0/1 witnesses (one bucket holding half a window) -- for G1 and G2, through the
per-window path and the resident-table path.  Expected values from the exponent side (oracle/c)."""
import numpy as np
import pytest

from oracle import bn254 as C
from oracle import cpu as ocpu

from helpers import assert_limbs_equal, fr_from_ints, fr_rand, g1_gen_arr, g2_gen_arr

pytestmark = pytest.mark.gpu
R = C.R
N = (1 << 17) + 123


@pytest.fixture()
def ba(ctx):
    yield ctx
    ctx.set("msm_ba_rounds", 0)


def _case(kind):
    k = fr_rand(8100, N)
    s = fr_rand(8101, N)
    if kind == "equal_bases":                 # every base the same point: every pair addition is a doubling
        k[:] = k[0]
    elif kind == "few_bases":                 # 7 distinct points: doublings and additions mixed in every bucket
        k[:] = k[np.arange(N) % 7]
    elif kind == "negatives":                 # P, -P alternating with equal scalars: pairs cancel to infinity
        ki = ocpu.fr_ints(k[: N // 2 + 1])
        both = np.empty_like(k)
        both[0::2] = fr_from_ints(ki)[: len(both[0::2])]
        both[1::2] = fr_from_ints([R - v for v in ki])[: len(both[1::2])]
        k = both
        s[1::2] = s[0::2][: len(s[1::2])]
    elif kind == "bits":                      # witness-like 0 / 1 / r-1 scalars: a few enormous buckets
        s[:] = 0
        s[1::2] = fr_from_ints([1])[0]
        s[5::64] = fr_from_ints([R - 1])[0]
    elif kind == "zero_and_inf":              # zero scalars and bases at infinity ([0]G) sprinkled in
        s[3::5] = 0
        k[7::11] = 0
    return k, s


@pytest.mark.parametrize("kind", ["random", "equal_bases", "few_bases", "negatives", "bits", "zero_and_inf"])
@pytest.mark.parametrize("rounds", [1, 3, 5])
def test_g1_batch_affine_rounds_match_trapdoor(ba, kind, rounds):
    k, s = _case(kind)
    bases = ba.g1_fixed_base_mul(k)
    want = ocpu.g1_mul(g1_gen_arr(), ocpu.fr_dot(s, k))
    ba.set("msm_ba_rounds", rounds)
    assert_limbs_equal(ba.msm_g1(s, bases), want, f"per-window path, {kind}, {rounds} rounds")
    h = ba.bases_upload(bases)
    try:
        assert_limbs_equal(ba.msm_resident(h, s), want, f"resident tables, {kind}, {rounds} rounds")
    finally:
        ba.bases_free(h)


@pytest.mark.parametrize("kind", ["random", "few_bases", "negatives", "bits"])
def test_g2_batch_affine_rounds_match_trapdoor(ba, kind):
    k, s = _case(kind)
    bases = ba.g2_fixed_base_mul(k)
    want = ocpu.g2_mul(g2_gen_arr(), ocpu.fr_dot(s, k))
    ba.set("msm_ba_rounds", 3)
    assert_limbs_equal(ba.msm_g2(s, bases), want, f"G2 per-window path, {kind}")
    h = ba.bases_upload(bases, g2=True)
    try:
        assert_limbs_equal(ba.msm_resident(h, s, g2=True), want, f"G2 resident tables, {kind}")
    finally:
        ba.bases_free(h)
