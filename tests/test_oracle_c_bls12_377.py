"""The BLS12-377 build of the C oracle (oracle/c, ORC_CURVE=377) against the Python oracle instance oracle.bls12_377
(the independent big-int implementation): constants, fields, both groups, the transforms, computeH, the bucket-method
MSMs and the pow17 MiMC-shaped chain with its trapdoor setup scalars.  CPU only; mirrors the BN254 tests in
tests/test_oracle.py."""
import numpy as np

from oracle.bls12_377 import cpu as ocpu
from oracle.bls12_377 import curve as C
from oracle.bls12_377 import groth16 as og
from oracle.bls12_377 import msm as omsm
from oracle.bls12_377 import ntt as ontt
from oracle.bls12_377 import r1cs as orcs
from oracle.prng import Xoshiro, SEED_BASE

R = C.R


def test_the_two_bindings_are_distinct_instances():
    from oracle import cpu as ocpu254
    assert (ocpu.NAME, ocpu254.NAME) == ("bls12_377", "bn254")
    assert (ocpu.G1W, ocpu.G2W, ocpu254.G1W, ocpu254.G2W) == (12, 24, 8, 16)
    assert ocpu.lib() is not ocpu254.lib()


def test_field_and_curve_vs_python():
    a, b = ocpu.fr_random(11, 50), ocpu.fr_random(12, 50)
    ai, bi = ocpu.fr_ints(a), ocpu.fr_ints(b)
    assert ocpu.fr_ints(ocpu.fr_mul_vec(a, b)) == [x * y % R for x, y in zip(ai, bi)]
    assert ocpu.fr_ints(ocpu.fr_dot(a, b).reshape(1, 4))[0] == sum(x * y for x, y in zip(ai, bi)) % R
    pts = ocpu.g1_mul_gen_batch(a[:5])
    assert pts.shape == (5, 12)
    assert [ocpu.g1_pt(p) for p in pts] == [C.g1_mul(C.G1_GEN, k) for k in ai[:5]]
    p2 = ocpu.g2_mul_gen_batch(a[:3])
    assert p2.shape == (3, 24)
    assert [ocpu.g2_pt(p) for p in p2] == [C.g2_mul(C.G2_GEN, k) for k in ai[:3]]
    # double-and-add, mixed addition (incl. doubling and the inverse pair) in both groups
    assert ocpu.g1_pt(ocpu.g1_mul(pts[0], b[0])) == C.g1_mul(C.G1_GEN, ai[0] * bi[0] % R)
    assert ocpu.g2_pt(ocpu.g2_mul(p2[0], b[0])) == C.g2_mul(C.G2_GEN, ai[0] * bi[0] % R)
    assert ocpu.g1_pt(ocpu.g1_add(pts[0], pts[1])) == C.g1_mul(C.G1_GEN, (ai[0] + ai[1]) % R)
    assert ocpu.g1_pt(ocpu.g1_add(pts[0], pts[0])) == C.g1_mul(C.G1_GEN, 2 * ai[0] % R)
    assert ocpu.g2_pt(ocpu.g2_add(p2[0], p2[1])) == C.g2_mul(C.G2_GEN, (ai[0] + ai[1]) % R)
    neg = ocpu.g1_mul_gen_batch(ocpu.fr_arr([(-ai[0]) % R]))[0]
    assert not ocpu.g1_add(pts[0], neg).any()
    # PRNG parity between the two oracles
    rng = Xoshiro(11)
    assert ai[:4] == [rng.field(R) for _ in range(4)]


def test_msm_vs_python_pippenger_and_trapdoor():
    n = 300
    k, s = ocpu.fr_random(21, n), ocpu.fr_random(22, n)
    ki, si = ocpu.fr_ints(k), ocpu.fr_ints(s)
    p1, p2 = ocpu.g1_mul_gen_batch(k), ocpu.g2_mul_gen_batch(k[:40])
    d = sum(x * y for x, y in zip(ki, si)) % R
    want1 = C.g1_mul(C.G1_GEN, d)
    assert ocpu.g1_pt(ocpu.msm_g1(p1, s)) == want1
    assert ocpu.g1_pt(ocpu.msm_g1(p1, s, c=7, nthreads=3)) == want1
    assert omsm.msm_pippenger([ocpu.g1_pt(p) for p in p1[:20]], si[:20], c=5) == ocpu.g1_pt(ocpu.msm_g1(p1[:20], s[:20]))
    d2 = sum(x * y for x, y in zip(ki[:40], si[:40])) % R
    assert ocpu.g2_pt(ocpu.msm_g2(p2, s[:40])) == C.g2_mul(C.G2_GEN, d2)
    assert omsm.msm_pippenger([ocpu.g2_pt(p) for p in p2[:6]], si[:6], c=4, g2=True) == ocpu.g2_pt(ocpu.msm_g2(p2[:6], s[:6]))
    assert not ocpu.msm_g1(p1[:0], s[:0]).any()


def test_transforms_and_compute_h_vs_python():
    n = 64
    x = ocpu.fr_random(31, n)
    xi = ocpu.fr_ints(x)
    d = ontt.Domain(n)
    assert ocpu.fr_ints(ocpu.fft(x, ocpu.FFT_INV_DIF)) == ontt.fft_inverse_dif(xi, d)
    assert ocpu.fr_ints(ocpu.fft(x, ocpu.FFT_DIT_COSET)) == ontt.fft_dit(xi, d, coset=True)
    assert ocpu.fr_ints(ocpu.fft(x, ocpu.FFT_INV_DIF_COSET)) == ontt.fft_inverse_dif(xi, d, coset=True)
    assert ocpu.fr_ints(ocpu.fft(x, ocpu.FFT_DIT)) == ontt.fft_dit(xi, d)
    nc = 50
    a, b = ocpu.fr_random(32, nc), ocpu.fr_random(33, nc)
    c = ocpu.fr_mul_vec(a, b)
    h = ocpu.compute_h(a, b, c, n)
    assert ocpu.fr_ints(h) == ontt.compute_h(ocpu.fr_ints(a), ocpu.fr_ints(b), ocpu.fr_ints(c), d)


def test_chain_matches_python_pow17_chain():
    """the pow17 MiMC-shaped chain generator + trapdoor setup scalars (C) against the Python R1CS path."""
    nc, seed = 43, SEED_BASE + 77
    rng = Xoshiro(seed)
    cs, consts = orcs.mimc_chain(nc, rng, power=17)
    assert len(consts) == (nc - 1) // 5 and ocpu.fr_ints(ocpu.mimc_chain_constants(nc, seed)) == consts
    x0, k = 1234567, 7654321
    w = orcs.mimc_chain_witness(cs, x0, k)
    L, _, _ = cs.constraints[-1]
    w[1] = sum(cf * w[i] for cf, i in L) % R
    Wc, a, b, c = ocpu.mimc_chain_witness(nc, seed, ocpu.fr_arr([x0])[0], ocpu.fr_arr([k])[0])
    assert ocpu.fr_ints(Wc) == w
    pa, pb, pc = cs.abc(w)
    assert (ocpu.fr_ints(a), ocpu.fr_ints(b), ocpu.fr_ints(c)) == (pa, pb, pc)
    tw = og.ToxicWaste(Xoshiro(3))
    d = ontt.Domain(nc)
    A, B, Cc = og.setup_abc(cs, d, tw)
    twa = ocpu.fr_arr([tw.t, tw.alpha, tw.beta, tw.gamma, tw.delta])
    Ac, Bc, Kc, Zc = ocpu.mimc_chain_setup_scalars(nc, d.n, seed, twa)
    assert ocpu.fr_ints(Ac) == A and ocpu.fr_ints(Bc) == B
    K = [(A[i] * tw.beta + B[i] * tw.alpha + Cc[i]) % R * (tw.gamma_inv if i < 2 else tw.delta_inv) % R for i in range(cs.nb_wires)]
    assert ocpu.fr_ints(Kc) == K
    zt = (pow(tw.t, d.n, R) - 1) * tw.delta_inv % R
    assert ocpu.fr_ints(Zc) == [zt * pow(tw.t, i, R) % R for i in range(d.n)]
    # the infinity pattern bench.py's chain_shape(power=17) assumes: the t^16 wire of every round is absent from R
    inf_b = np.array([v == 0 for v in B])
    assert inf_b[1] and inf_b[7::5][: len(consts)].all()
