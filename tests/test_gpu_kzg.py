"""GPU: the other consumers of the MSM kernels -- kzg.Commit as the PLONK prover calls it
(backend/plonk/bn254/prove.go:312,460,718,1158-1172) and the MPC-setup linear combinations
(backend/groth16/bn254/mpcsetup/utils.go:120-154) -- against the oracle: a commitment to p under the SRS [tau^i]G must
be [p(tau)]G (the trapdoor side of KZG), and equal the oracle's own MultiExp."""
import numpy as np
import pytest

from gnark_icicle_b200 import kzg
from oracle import bn254 as C
from oracle import cpu as ocpu

from helpers import assert_limbs_equal, fr_from_ints, fr_rand, g1_bases, g2_bases, g1_gen_arr

pytestmark = pytest.mark.gpu
R = C.R


@pytest.fixture(scope="module")
def srs(ctx):
    tau, n = 0x1F2E3D4C5B6A79881726354453627180 % R, 1 << 15
    pw, t = [], 1
    for _ in range(n):
        pw.append(t)
        t = t * tau % R
    key = kzg.ProvingKey(G1=ctx.g1_fixed_base_mul(fr_from_ints(pw)))
    yield tau, pw, key
    key.free_device()


@pytest.mark.parametrize("deg", [1, 7, 5000, (1 << 15) - 1, 1 << 15])
def test_kzg_commit_is_p_of_tau_times_g(ctx, srs, deg):
    tau, pw, key = srs
    p = fr_rand(300 + deg % 97, deg)
    got = kzg.Commit(p, key, ctx)
    p_tau = ocpu.fr_dot(p, fr_from_ints(pw[:deg]))                 # p(tau) in Fr (Montgomery limbs)
    assert_limbs_equal(got, ocpu.g1_mul(g1_gen_arr(), p_tau), f"kzg.Commit of a degree-{deg - 1} polynomial")
    assert_limbs_equal(got, ocpu.msm_g1(key.G1[:deg], p), "kzg.Commit vs the oracle's MultiExp over pk.G1[:len(p)]")


def test_kzg_commit_rejects_bad_sizes_and_commits_quotient_shards(ctx, srs):
    tau, pw, key = srs
    with pytest.raises(kzg.ErrInvalidPolynomialSize):
        kzg.Commit(np.zeros((0, 4), dtype=np.uint64), key, ctx)
    with pytest.raises(kzg.ErrInvalidPolynomialSize):
        kzg.Commit(fr_rand(1, (1 << 15) + 1), key, ctx)
    hs = [fr_rand(400 + i, 3000 + i) for i in range(3)]
    for h, d in zip(hs, kzg.commit_to_quotient(*hs, key, ctx)):
        assert_limbs_equal(d, ocpu.msm_g1(key.G1[:len(h)], h), "commitToQuotient shard")


def test_mpcsetup_linear_combinations(ctx):
    n = 3000
    _, A = g1_bases(501, n)
    _, B = g1_bases(502, n)
    r = fr_rand(503, n)
    a, b = kzg.merge(A, B, r, ctx)
    assert_limbs_equal(a, ocpu.msm_g1(A, r), "merge: sum r_i A_i")
    assert_limbs_equal(b, ocpu.msm_g1(B, r), "merge: sum r_i B_i")
    L1, L2 = kzg.linear_combination_g1(A, r[:-1], ctx)
    assert_limbs_equal(L1, ocpu.msm_g1(A[:-1], r[:-1]), "linearCombinationG1 L1")
    assert_limbs_equal(L2, ocpu.msm_g1(A[1:], r[:-1]), "linearCombinationG1 L2")
    _, A2 = g2_bases(504, 700)
    L1, L2 = kzg.linear_combination_g2(A2, r[:699], ctx)
    assert_limbs_equal(L1, ocpu.msm_g2(A2[:-1], r[:699]), "linearCombinationG2 L1")
    assert_limbs_equal(L2, ocpu.msm_g2(A2[1:], r[:699]), "linearCombinationG2 L2")
