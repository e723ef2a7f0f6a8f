"""BLS12-377 (BASELINE config 5) parity at sizes the Python oracle cannot reach, through the g16_bls12_377_* C ABI
against the BLS12-377 build of the C oracle (oracle/c with ORC_CURVE=377, itself checked against the Python oracle in
tests/test_oracle_c_bls12_377.py): bucket-method MSMs in both groups, the transforms and computeH, and the five points
of a prove on the pow17 MiMC-shaped chain -- against the CPU prover AND the trapdoor identity.  Bit-exact throughout."""
import numpy as np
import pytest

from gnark_icicle_b200 import lib
from oracle.bls12_377 import cpu as ocpu

from helpers import ChainCase, assert_limbs_equal

pytestmark = pytest.mark.gpu
CURVE = "bls12_377"


@pytest.fixture(scope="module")
def ctx377():
    c = lib.Context(0, curve=CURVE)
    yield c
    c.close()


@pytest.mark.parametrize("logn", [16, 20])
def test_msm_g1_matches_c_oracle(ctx377, logn):
    n = (1 << logn) - 5
    k, s = ocpu.fr_random(7000 + logn, n), ocpu.fr_random(7100 + logn, n)
    bases = ctx377.g1_fixed_base_mul(k)
    assert_limbs_equal(bases[:300], ocpu.g1_mul_gen_batch(k[:300]), "fixed-base G1")
    got = ctx377.msm_g1(s, bases)
    assert_limbs_equal(got, ocpu.msm_g1(bases, s), f"G1 MSM 2^{logn} vs the oracle's bucket method")
    assert_limbs_equal(got, ocpu.g1_mul(ocpu.g1_gen_arr(), ocpu.fr_dot(s, k)), f"G1 MSM 2^{logn} vs the trapdoor")


@pytest.mark.parametrize("logn", [14, 18])
def test_msm_g2_matches_c_oracle(ctx377, logn):
    n = (1 << logn) + 3
    k, s = ocpu.fr_random(7200 + logn, n), ocpu.fr_random(7300 + logn, n)
    bases = ctx377.g2_fixed_base_mul(k)
    assert_limbs_equal(bases[:100], ocpu.g2_mul_gen_batch(k[:100]), "fixed-base G2")
    got = ctx377.msm_g2(s, bases)
    assert_limbs_equal(got, ocpu.msm_g2(bases, s), f"G2 MSM 2^{logn} vs the oracle's bucket method")
    assert_limbs_equal(got, ocpu.g2_mul(ocpu.g2_gen_arr(), ocpu.fr_dot(s, k)), f"G2 MSM 2^{logn} vs the trapdoor")


@pytest.mark.parametrize("logn", [18, 22])
def test_transforms_and_compute_h_match_c_oracle(ctx377, logn):
    n = 1 << logn
    x = ocpu.fr_random(7400 + logn, n)
    assert_limbs_equal(ctx377.ntt(x, inverse=True, coset=False, decimation=lib.DIF), ocpu.fft(x, ocpu.FFT_INV_DIF), "FFTInverse(DIF)")
    assert_limbs_equal(ctx377.ntt(x, inverse=False, coset=True, decimation=lib.DIT), ocpu.fft(x, ocpu.FFT_DIT_COSET), "FFT(DIT, OnCoset)")
    assert_limbs_equal(ctx377.ntt(x, inverse=True, coset=True, decimation=lib.DIF), ocpu.fft(x, ocpu.FFT_INV_DIF_COSET), "FFTInverse(DIF, OnCoset)")
    nc = n - 7
    a, b = x[:nc], ocpu.fr_random(7500 + logn, nc)
    c = ocpu.fr_mul_vec(a, b)
    c[5] = a[5]                      # not a satisfied system: computeH is still a well-defined map
    assert_limbs_equal(ctx377.compute_h(a, b, c, n), ocpu.compute_h(a, b, c, n), f"computeH 2^{logn}")


@pytest.mark.parametrize("logn", [16, 20])
def test_prove_pow17_chain_matches_cpu_prover_and_trapdoor(ctx377, logn):
    """config 5's shape: the five points of g16_bls12_377_prove against oracle/c's CPU prover (computeH + five
    bucket-method MultiExps on the same key) and against the exponent side."""
    nc = 1 << logn
    case = ChainCase(nc, seed=377, g1_mul=ctx377.g1_fixed_base_mul, g2_mul=ctx377.g2_fixed_base_mul, oc=ocpu)
    pk = case.pk
    assert pk.g1_a.shape[1] == 12 and pk.g2_b.shape[1] == 24
    # the key the GPU built is the key the oracle would have built
    assert_limbs_equal(pk.g1_k[:200], ocpu.g1_mul_gen_batch(case.sK[:200]), "key G1.K")
    handle = ctx377.pk_upload(pk.domain_n, pk.g1_a, pk.g1_b, pk.g1_k, pk.g1_z, pk.g2_b, pk.infinity_a, pk.infinity_b, pk.nb_wires)
    try:
        pts, st = ctx377.prove(handle, case.W, case.a, case.b, case.c)
        pts2, _ = ctx377.prove(handle, case.W, case.a, case.b, case.c)
    finally:
        ctx377.pk_free(handle)
    h = ocpu.compute_h(case.a, case.b, case.c, case.n)
    wa, wb, wk = case.W[~case.inf_a], case.W[~case.inf_b], case.W[2:]
    cpu = np.concatenate([ocpu.msm_g1(pk.g1_a, wa), ocpu.msm_g1(pk.g1_b, wb), ocpu.msm_g1(pk.g1_k, wk),
                          ocpu.msm_g1(pk.g1_z, h[: case.n - 1]), ocpu.msm_g2(pk.g2_b, wb)])
    assert_limbs_equal(pts, cpu, f"2^{logn} prove points vs the CPU prover")
    assert_limbs_equal(pts, case.expected_points(h), f"2^{logn} prove points vs the trapdoor identity")
    assert_limbs_equal(pts2, pts, "second prove on the same key")
    assert st["msm_n"][3] == case.n - 1
