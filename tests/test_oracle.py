"""CPU-only: pins the oracle (the checker) -- reference KATs, algebraic self-validation, golden fixtures,
and C oracle == Python oracle.  Nothing here touches the product library."""
import json
import os

import numpy as np
import pytest

from oracle import bn254 as C
from oracle import cpu as ocpu
from oracle import groth16 as og
from oracle import msm as omsm
from oracle import ntt as ontt
from oracle import r1cs as orcs
from oracle.prng import Xoshiro, SEED_BASE

R = C.R
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def gold(name):
    with open(os.path.join(GOLD, name + ".json")) as f:
        return json.load(f)


def limbs(hexstr, width):
    return np.frombuffer(bytes.fromhex(hexstr), dtype=np.uint64).reshape(-1, width).copy()


# ---- reference KATs (backend/groth16/bn254/utils_test.go:17-38)
def test_filter_heap_kats():
    elems = [0, 1, 2, 3]
    assert og.filter_heap(elems, 0, [1, 2]) == [0, 3]
    assert og.filter_heap(elems[1:], 1, [1, 2]) == [3]
    assert og.filter_heap(elems, 0, [1, 1, 2]) == [0, 3]
    assert og.filter_heap(elems[1:], 1, [1, 1, 2]) == [3]


# ---- examples/cubic/cubic_test.go:23-38: (3, 35) proves and verifies, (42, 42) must fail
def test_cubic_valid_and_invalid():
    cs = orcs.cubic_circuit()
    rng = Xoshiro(SEED_BASE + 1)
    pk, vk, _ = og.setup(cs, rng)
    w = orcs.cubic_witness(3, 35)
    proof = og.prove(cs, pk, w, rng.field(R), rng.field(R))
    assert og.verify(proof, vk, [35])
    assert not og.verify(proof, vk, [34])
    with pytest.raises(ValueError):
        cs.abc(orcs.cubic_witness(42, 42))


def test_cubic_proof_matches_golden_and_exponent_side():
    g = gold("cubic_proof")
    cs = orcs.cubic_circuit()
    rng = Xoshiro(SEED_BASE + 1)
    pk, vk, scal = og.setup(cs, rng)
    r, s = rng.field(R), rng.field(R)
    w = orcs.cubic_witness(3, 35)
    proof = og.prove(cs, pk, w, r, s)
    assert og.proof_bytes(proof).hex() == g["proof"]
    # exponent side: Ar = [alpha + sum w_i A_i(t) + r delta] G with the known toxic waste
    tw = scal["tw"]
    wa = [w[i] for i in range(len(w)) if not pk.infinity_a[i]]
    e = (tw.alpha + sum(x * y for x, y in zip(wa, scal["A"])) + r * tw.delta) % R
    assert proof["ar"] == C.g1_mul(C.G1_GEN, e)


# ---- NTT: against O(n^2) evaluation, and the quotient identity
def test_ntt_vs_naive_evaluation():
    n = 16
    d = ontt.Domain(n)
    rng = Xoshiro(5)
    coeffs = [rng.field(R) for _ in range(n)]
    br = ontt.bit_reverse_list(coeffs)
    assert ontt.fft_dit(br, d) == [ontt.poly_eval(coeffs, pow(d.gen, k, R)) for k in range(n)]
    assert ontt.fft_dit(br, d, coset=True) == [ontt.poly_eval(coeffs, C.FR_GEN * pow(d.gen, k, R) % R) for k in range(n)]
    ev = [ontt.poly_eval(coeffs, pow(d.gen, k, R)) for k in range(n)]
    assert ontt.bit_reverse_list(ontt.fft_inverse_dif(ev, d)) == coeffs


def test_compute_h_quotient_identity():
    n = 32
    d = ontt.Domain(n)
    rng = Xoshiro(6)
    a = [rng.field(R) for _ in range(n)]
    b = [rng.field(R) for _ in range(n)]
    c = [x * y % R for x, y in zip(a, b)]
    h = ontt.bit_reverse_list(ontt.compute_h(a, b, c, d))
    z = rng.field(R)
    pa, pb, pc = (ontt.bit_reverse_list(ontt.fft_inverse_dif(v, d)) for v in (a, b, c))
    lhs = (ontt.poly_eval(pa, z) * ontt.poly_eval(pb, z) - ontt.poly_eval(pc, z)) % R
    assert lhs == ontt.poly_eval(h, z) * (pow(z, n, R) - 1) % R


# ---- MSM: bucket method vs double-and-add vs trapdoor
def test_pippenger_vs_naive_vs_trapdoor():
    rng = Xoshiro(7)
    n = 20
    ks = [rng.field(R) for _ in range(n)]
    sc = [rng.field(R) for _ in range(n)]
    pts = [C.g1_mul(C.G1_GEN, k) for k in ks]
    want = C.g1_mul(C.G1_GEN, sum(a * b for a, b in zip(ks, sc)) % R)
    assert omsm.msm_naive(pts, sc) == want
    assert omsm.msm_pippenger(pts, sc, c=5) == want
    p2 = [C.g2_mul(C.G2_GEN, k) for k in ks[:6]]
    want2 = C.g2_mul(C.G2_GEN, sum(a * b for a, b in zip(ks[:6], sc[:6])) % R)
    assert omsm.msm_pippenger(p2, sc[:6], c=4, g2=True) == want2


# ---- golden fixtures vs the C oracle (independent implementation)
def test_golden_msm_vs_c_oracle():
    g = gold("msm_small")
    s = limbs(g["scalars"], 4)
    assert np.array_equal(ocpu.msm_g1(limbs(g["g1_points"], 8), s), limbs(g["g1_result"], 8)[0])
    assert np.array_equal(ocpu.msm_g2(limbs(g["g2_points"], 16), s), limbs(g["g2_result"], 16)[0])
    assert np.array_equal(ocpu.msm_g1(limbs(g["g1_points"], 8), s, c=7, nthreads=3), limbs(g["g1_result"], 8)[0])


def test_golden_ntt_vs_c_oracle():
    g = gold("ntt_small")
    x = limbs(g["x"], 4)
    for mode, key in ((ocpu.FFT_INV_DIF, "inv_dif"), (ocpu.FFT_DIT_COSET, "dit_coset"),
                      (ocpu.FFT_INV_DIF_COSET, "inv_dif_coset"), (ocpu.FFT_DIT, "dit")):
        assert np.array_equal(ocpu.fft(x, mode), limbs(g[key], 4)), key
    h = ocpu.compute_h(limbs(g["h_a"], 4), limbs(g["h_b"], 4), limbs(g["h_c"], 4), g["n"])
    assert np.array_equal(h, limbs(g["h"], 4))


def test_c_oracle_field_and_curve_vs_python():
    a, b = ocpu.fr_random(11, 50), ocpu.fr_random(12, 50)
    ai, bi = ocpu.fr_ints(a), ocpu.fr_ints(b)
    assert ocpu.fr_ints(ocpu.fr_mul_vec(a, b)) == [x * y % R for x, y in zip(ai, bi)]
    assert ocpu.fr_ints(ocpu.fr_dot(a, b).reshape(1, 4))[0] == sum(x * y for x, y in zip(ai, bi)) % R
    pts = ocpu.g1_mul_gen_batch(a[:5])
    assert [ocpu.g1_pt(p) for p in pts] == [C.g1_mul(C.G1_GEN, k) for k in ai[:5]]
    p2 = ocpu.g2_mul_gen_batch(a[:3])
    assert [ocpu.g2_pt(p) for p in p2] == [C.g2_mul(C.G2_GEN, k) for k in ai[:3]]
    # PRNG parity between the two oracles
    rng = Xoshiro(11)
    assert ai[:4] == [rng.field(R) for _ in range(4)]


def test_c_oracle_chain_matches_python_chain():
    """the MiMC-shaped chain generator + trapdoor setup scalars (C) against the Python R1CS path."""
    nc, seed = 40, SEED_BASE + 77
    rng = Xoshiro(seed)
    cs, _ = orcs.mimc_chain(nc, rng)
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
