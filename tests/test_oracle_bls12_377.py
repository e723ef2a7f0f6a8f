"""CPU-only, BLS12-377 (BASELINE config 5): pins the second oracle instance (constants, NTT, MSM, Groth16 algebra) and
checks the library's BLS12-377 entry points exist and its host-side group helpers (12-limb Fp, Fp2 over u^2 = -5)
agree with the oracle.  No golden vectors exist for this curve in the reference (it has no BLS12-377 icicle
package): parity unpinned, self-validated algebraically like the BN254 oracle."""
import os
import re

import numpy as np

from gnark_icicle_b200 import lib
from oracle.bls12_377 import curve as C
from oracle.bls12_377 import groth16 as og
from oracle.bls12_377 import msm as omsm
from oracle.bls12_377 import ntt as ontt
from oracle.bls12_377 import r1cs as orcs
from oracle.prng import Xoshiro, SEED_BASE

from helpers377 import fr_arr, g1_arr, g2_arr, g1_pts, g2_pts

R, P = C.R, C.P
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- constants
def test_curve_parameters_follow_from_the_seed():
    x = 0x8508C00000000001
    assert R == x ** 4 - x ** 2 + 1
    assert P == ((x - 1) ** 2 * R) // 3 + x and ((x - 1) ** 2 * R) % 3 == 0
    assert C.NAME == "bls12_377" and C.FP_BYTES == 48
    assert (R - 1) % (1 << 47) == 0 and (R - 1) % (1 << 48) != 0
    # gnark-crypto's published 2^47-th root of unity of bls12-377 fr
    assert C.FR_ROOT == 8065159656716812877374967518403273466521432693661810619979959746626482506078
    assert pow(C.FR_ROOT, 1 << 47, R) == 1 and pow(C.FR_ROOT, 1 << 46, R) == R - 1
    assert pow(C.FR_GEN, (R - 1) // 2, R) == R - 1            # 22 generates the 2-part: a valid coset shift
    assert pow(-5 % P, (P - 1) // 2, P) == P - 1              # -5 is a non-residue: Fp2 = Fp[u]/(u^2+5) is a field


def test_generators_are_on_curve_and_of_order_r():
    assert C.g1_on_curve(C.G1_GEN) and C.g1_mul(C.G1_GEN, R) is None
    assert C.g2_on_curve(C.G2_GEN) and C.g2_mul(C.G2_GEN, R) is None
    assert C.B2 == (0, (-pow(5, -1, P)) % P)                  # 1/u = -u/5
    # group law sanity: (a+b)G == aG + bG on both groups
    a, b = 0x1234567890ABCDEF1234567, R - 98765
    assert C.g1_mul(C.G1_GEN, (a + b) % R) == C.g1_add(C.g1_mul(C.G1_GEN, a), C.g1_mul(C.G1_GEN, b))
    assert C.g2_mul(C.G2_GEN, (a + b) % R) == C.g2_add(C.g2_mul(C.G2_GEN, a), C.g2_mul(C.G2_GEN, b))


def test_memory_image_round_trips():
    rng = Xoshiro(SEED_BASE + 50)
    v = [rng.field(R) for _ in range(5)]
    assert C.fr_vec_from_bytes(C.fr_vec_to_bytes(v)) == v
    p1, p2 = C.g1_mul(C.G1_GEN, v[0]), C.g2_mul(C.G2_GEN, v[1])
    assert len(C.g1_to_bytes(p1)) == 96 and len(C.g2_to_bytes(p2)) == 192
    assert C.g1_from_bytes(C.g1_to_bytes(p1)) == p1 and C.g2_from_bytes(C.g2_to_bytes(p2)) == p2
    assert C.g1_from_bytes(bytes(96)) is None and C.g2_to_bytes(None) == bytes(192)


# ---- NTT / computeH
def test_ntt_matches_polynomial_evaluation_and_quotient_identity():
    rng = Xoshiro(SEED_BASE + 51)
    n = 16
    d = ontt.Domain(n)
    coeffs = [rng.field(R) for _ in range(n)]
    ev = ontt.fft_dit(ontt.bit_reverse_list(coeffs), d)
    assert ev == [ontt.poly_eval(coeffs, pow(d.gen, i, R)) for i in range(n)]
    evc = ontt.fft_dit(ontt.bit_reverse_list(coeffs), d, coset=True)
    assert evc == [ontt.poly_eval(coeffs, C.FR_GEN * pow(d.gen, i, R) % R) for i in range(n)]
    back = ontt.bit_reverse_list(ontt.fft_inverse_dif(ev, d))
    assert back == coeffs
    # h (X^n - 1) == A B - C at a random point, for a, b, c with c = a o b on the domain
    a = [rng.field(R) for _ in range(n)]
    b = [rng.field(R) for _ in range(n)]
    c = [x * y % R for x, y in zip(a, b)]
    h = ontt.bit_reverse_list(ontt.compute_h(a, b, c, d))
    pa, pb, pc = (ontt.bit_reverse_list(ontt.fft_inverse_dif(v, d)) for v in (a, b, c))
    z = rng.field(R)
    lhs = ontt.poly_eval(h, z) * (pow(z, n, R) - 1) % R
    assert lhs == (ontt.poly_eval(pa, z) * ontt.poly_eval(pb, z) - ontt.poly_eval(pc, z)) % R


# ---- MSM
def test_pippenger_matches_naive_and_trapdoor():
    rng = Xoshiro(SEED_BASE + 52)
    n = 24
    ks = [rng.field(R) for _ in range(n)]
    ss = [rng.field(R) for _ in range(n)]
    ss[3], ss[4], ss[5] = 0, 1, R - 1
    p1 = omsm.batch_scalar_mul_gen(ks)
    p1[7] = None                                   # an infinity base is legal
    want = C.g1_mul(C.G1_GEN, sum(s * k for i, (s, k) in enumerate(zip(ss, ks)) if i != 7) % R)
    assert omsm.msm_naive(p1, ss) == want
    for c in (4, 7, 11):
        assert omsm.msm_pippenger(p1, ss, c=c) == want
    p2 = omsm.batch_scalar_mul_gen(ks[:8], g2=True)
    want2 = C.g2_mul(C.G2_GEN, sum(s * k for s, k in zip(ss, ks[:8])) % R)
    assert omsm.msm_pippenger(p2, ss[:8], c=5, g2=True) == want2


# ---- Groth16 algebra: exponent-side check with the toxic waste, then the restated pairing check
def test_proof_satisfies_the_exponent_side_equations():
    rng = Xoshiro(SEED_BASE + 53)
    cs, _ = orcs.mimc_chain(22, rng, power=17)
    pk, vk, scal = og.setup(cs, rng)
    w = orcs.mimc_chain_witness(cs, rng.field(R), rng.field(R))
    w[1] = (sum(cf * w[i] for cf, i in cs.constraints[-1][0]) * 1) % R   # the public output the last constraint binds
    a, b, c = cs.abc(w)
    r, s = rng.field(R), rng.field(R)
    proof = og.prove(cs, pk, w, r, s)
    tw = scal["tw"]
    wa = [w[i] for i in range(len(w)) if not pk.infinity_a[i]]
    wb = [w[i] for i in range(len(w)) if not pk.infinity_b[i]]
    A = (tw.alpha + sum(x * y for x, y in zip(wa, scal["A"])) + r * tw.delta) % R
    B = (tw.beta + sum(x * y for x, y in zip(wb, scal["B"])) + s * tw.delta) % R
    assert proof["ar"] == C.g1_mul(C.G1_GEN, A)
    assert proof["bs"] == C.g2_mul(C.G2_GEN, B)
    # Groth16: A B = alpha beta + (public part) gamma + C delta, all in the exponent
    h = ontt.compute_h(a, b, c, pk.domain)
    Cexp = (sum(x * y for x, y in zip(w[pk.nb_public:], scal["K"])) + sum(x * y for x, y in zip(h, scal["Z"]))
            + s * A + r * B - r * s * tw.delta) % R
    assert proof["krs"] == C.g1_mul(C.G1_GEN, Cexp)


def test_pairing_is_bilinear_and_nondegenerate():
    from oracle.bls12_377 import pairing as PR
    a, b = 0x1234567, 0x89abcdef1
    e = PR.pairing_product([(C.g1_mul(C.G1_GEN, a), C.g2_mul(C.G2_GEN, b))])
    base = PR.pairing_product([(C.G1_GEN, C.G2_GEN)])
    assert e == PR.f12_pow(base, a * b % R) and base != PR.F12_ONE and PR.f12_pow(base, R) == PR.F12_ONE
    assert PR.pairing_product([(C.g1_mul(C.G1_GEN, a), C.G2_GEN), (C.g1_neg(C.G1_GEN), C.g2_mul(C.G2_GEN, a))]) == PR.F12_ONE


def test_proofs_verify_under_the_pairing_check():
    """verify.go:49-148 restated over BLS12-377: the oracle prover's proofs are accepted, a wrong public input, a tampered
    point and a proof for another witness are not (cubic: examples/cubic/cubic_test.go:23-38; pow17 chain)."""
    rng = Xoshiro(SEED_BASE + 57)
    cs, w = orcs.cubic_circuit(), orcs.cubic_witness(3, 35)
    pk, vk, _ = og.setup(cs, rng)
    proof = og.prove(cs, pk, w, rng.field(R), rng.field(R))
    assert og.verify(proof, vk, [35])
    assert not og.verify(proof, vk, [36])
    bad = dict(proof)
    bad["krs"] = C.g1_add(proof["krs"], C.G1_GEN)
    assert not og.verify(bad, vk, [35])
    cs2, _ = orcs.mimc_chain(31, rng, power=17)
    pk2, vk2, _ = og.setup(cs2, rng)
    w2 = orcs.mimc_chain_witness(cs2, rng.field(R), rng.field(R))
    w2[1] = sum(cf * w2[i] for cf, i in cs2.constraints[-1][0]) % R
    proof2 = og.prove(cs2, pk2, w2, rng.field(R), rng.field(R))
    assert og.verify(proof2, vk2, [w2[1]])
    assert not og.verify(proof2, vk2, [(w2[1] + 1) % R])


# ---- the library: BLS12-377 entry points and host helpers
def test_library_exports_the_bls12_377_entry_points():
    with open(os.path.join(ROOT, "include", "g16b200_bls12_377.h")) as f:
        hdr = f.read()
    declared = set(re.findall(r"\b(g16_bls12_377_[a-z0-9_]+)\s*\(", hdr))
    assert declared == {lib.symbol_name(s, "bls12_377") for s in lib.SYMBOLS}, declared ^ {lib.symbol_name(s, "bls12_377") for s in lib.SYMBOLS}
    handle = lib.load("bls12_377")
    for s in lib.SYMBOLS:
        assert getattr(handle, s) is not None
    # the generated headers are current
    import subprocess, sys
    before = hdr
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "gen_curve_abi.py")], stdout=subprocess.DEVNULL)
    with open(os.path.join(ROOT, "include", "g16b200_bls12_377.h")) as f:
        assert f.read() == before, "include/g16b200_bls12_377.h is stale: run tools/gen_curve_abi.py"


def test_host_group_helpers_match_oracle():
    rng = Xoshiro(SEED_BASE + 54)
    k = [rng.field(R) for _ in range(4)]
    g1, g2 = g1_arr([C.G1_GEN])[0], g2_arr([C.G2_GEN])[0]
    p = lib.g1_scalar_mul(g1, fr_arr([k[0]])[0], curve="bls12_377")
    assert g1_pts(p) == [C.g1_mul(C.G1_GEN, k[0])]
    q = lib.g2_scalar_mul(g2, fr_arr([k[1]])[0], curve="bls12_377")
    assert g2_pts(q) == [C.g2_mul(C.G2_GEN, k[1])]
    assert g1_pts(lib.g1_add(p, g1, curve="bls12_377")) == [C.g1_mul(C.G1_GEN, (k[0] + 1) % R)]
    assert g2_pts(lib.g2_add(q, g2, curve="bls12_377")) == [C.g2_mul(C.G2_GEN, (k[1] + 1) % R)]
    # identities: P + inf, P + P (doubling path), P + (-P)
    z1, z2 = np.zeros(12, dtype=np.uint64), np.zeros(24, dtype=np.uint64)
    assert np.array_equal(lib.g1_add(p, z1, curve="bls12_377"), p) and np.array_equal(lib.g2_add(z2, q, curve="bls12_377"), q)
    assert g1_pts(lib.g1_add(p, p, curve="bls12_377")) == [C.g1_mul(C.G1_GEN, 2 * k[0] % R)]
    assert g2_pts(lib.g2_add(q, q, curve="bls12_377")) == [C.g2_mul(C.G2_GEN, 2 * k[1] % R)]
    neg = g1_arr([C.g1_neg(C.g1_mul(C.G1_GEN, k[0]))])[0]
    assert not lib.g1_add(p, neg, curve="bls12_377").any()


def test_host_epilogue_matches_oracle_assemble():
    rng = Xoshiro(SEED_BASE + 55)
    cs = orcs.cubic_circuit()
    pk, vk, _ = og.setup(cs, rng)
    w = orcs.cubic_witness(3, 35)
    r, s = rng.field(R), rng.field(R)
    parts = og.prove_parts(cs, pk, w)
    want = og.assemble(pk, parts, r, s)
    pk_points = np.concatenate([g1_arr([pk.g1_alpha, pk.g1_beta, pk.g1_delta]).reshape(-1), g2_arr([pk.g2_beta, pk.g2_delta]).reshape(-1)])
    msm_points = np.concatenate([g1_arr([parts["ar"], parts["bs1"], parts["krs_k"], parts["krs_hz"]]).reshape(-1), g2_arr([parts["bs2"]]).reshape(-1)])
    got = lib.assemble_proof(pk_points, fr_arr([r])[0], fr_arr([s])[0], msm_points, curve="bls12_377")
    assert got.tobytes() == og.proof_bytes(want)
