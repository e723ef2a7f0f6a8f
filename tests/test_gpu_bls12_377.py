"""GPU parity tests for the BLS12-377 build (BASELINE config 5) through the g16_bls12_377_* C ABI, against the Python
oracle instance oracle.bls12_377 (bit-exact) and, at sizes the oracle cannot reach, the trapdoor identity
sum s_i [k_i]G = [sum s_i k_i]G with bases produced by the fixed-base kernel (itself checked against the oracle)."""
import numpy as np
import pytest

from gnark_icicle_b200 import lib
import gnark_icicle_b200 as g16
from oracle.bls12_377 import curve as C
from oracle.bls12_377 import groth16 as og
from oracle.bls12_377 import msm as omsm
from oracle.bls12_377 import ntt as ontt
from oracle.bls12_377 import r1cs as orcs
from oracle.prng import Xoshiro, SEED_BASE

from helpers377 import fr_arr, fr_ints, g1_arr, g2_arr, g1_pts, g2_pts, fr_rand, dot

pytestmark = pytest.mark.gpu
R = C.R
CURVE = "bls12_377"


@pytest.fixture(scope="module")
def ctx377():
    c = lib.Context(0, curve=CURVE)
    yield c
    c.close()


def test_wrong_curve_id_is_refused():
    import ctypes
    h = ctypes.c_void_p()
    ids = (ctypes.c_int * 1)(0)
    assert lib.load(CURVE).g16_ctx_create(ctypes.byref(h), 0, ids, 1) == 3      # BN254 id on the BLS12-377 entry point
    assert lib.load("bn254").g16_ctx_create(ctypes.byref(h), 1, ids, 1) == 3


# ---------------------------------------------------------------- field
def test_fr_mul_vec(ctx377):
    rng = Xoshiro(SEED_BASE + 60)
    a, b = fr_rand(rng, 777), fr_rand(rng, 777)
    a[:4] = [0, 1, R - 1, R - 1]
    b[:4] = [5, R - 1, R - 1, 1]
    got = ctx377.fr_mul_vec(fr_arr(a), fr_arr(b))
    assert fr_ints(got) == [x * y % R for x, y in zip(a, b)]


# ---------------------------------------------------------------- NTT / computeH
@pytest.mark.parametrize("logn", [0, 1, 3, 7, 10, 13])
def test_ntt_variants_match_oracle(ctx377, logn):
    rng = Xoshiro(SEED_BASE + 61 + logn)
    n = 1 << logn
    d = ontt.Domain(n)
    x = fr_rand(rng, n)
    for coset in (False, True):
        got = fr_ints(ctx377.ntt(fr_arr(x), inverse=True, coset=coset, decimation=lib.DIF))
        assert got == ontt.fft_inverse_dif(list(x), d, coset=coset), ("inverse DIF", coset)
        got = fr_ints(ctx377.ntt(fr_arr(x), inverse=False, coset=coset, decimation=lib.DIT))
        assert got == ontt.fft_dit(list(x), d, coset=coset), ("forward DIT", coset)


@pytest.mark.parametrize("nc", [3, 1000, 1 << 14])
def test_compute_h_matches_oracle(ctx377, nc):
    rng = Xoshiro(SEED_BASE + 62)
    n = 1
    while n < nc:
        n <<= 1
    n = max(n, 2)
    a, b = fr_rand(rng, nc), fr_rand(rng, nc)
    c = [x * y % R for x, y in zip(a, b)]
    for i in range(0, nc, 3):
        c[i] = rng.field(R)              # parity does not need divisibility
    pad = [0] * (n - nc)
    want = ontt.compute_h(a + pad, b + pad, c + pad, ontt.Domain(n))
    got = fr_ints(ctx377.compute_h(fr_arr(a), fr_arr(b), fr_arr(c), n))
    assert got == want


# ---------------------------------------------------------------- fixed base + MSM
def test_fixed_base_matches_oracle(ctx377):
    rng = Xoshiro(SEED_BASE + 63)
    k = [0, 1, 2, R - 1, 255, 256, 1 << 248] + fr_rand(rng, 30)
    assert g1_pts(ctx377.g1_fixed_base_mul(fr_arr(k))) == [C.g1_mul(C.G1_GEN, v) for v in k]
    k2 = k[:12]
    assert g2_pts(ctx377.g2_fixed_base_mul(fr_arr(k2))) == [C.g2_mul(C.G2_GEN, v) for v in k2]


@pytest.mark.parametrize("n", [0, 1, 2, 33, 700])
def test_msm_g1_matches_oracle_pippenger(ctx377, n):
    rng = Xoshiro(SEED_BASE + 64 + n)
    k, s = fr_rand(rng, n), fr_rand(rng, n)
    bases = ctx377.g1_fixed_base_mul(fr_arr(k)) if n else np.zeros((0, 12), dtype=np.uint64)
    got = g1_pts(ctx377.msm_g1(fr_arr(s), bases))[0]
    assert got == C.g1_mul(C.G1_GEN, dot(s, k))
    if 0 < n <= 33:
        assert got == omsm.msm_pippenger(g1_pts(bases), s, c=5)


@pytest.mark.parametrize("n", [1, 50, 3000])
def test_msm_g2_matches_trapdoor(ctx377, n):
    rng = Xoshiro(SEED_BASE + 65 + n)
    k, s = fr_rand(rng, n), fr_rand(rng, n)
    bases = ctx377.g2_fixed_base_mul(fr_arr(k))
    got = g2_pts(ctx377.msm_g2(fr_arr(s), bases))[0]
    assert got == C.g2_mul(C.G2_GEN, dot(s, k))
    if n == 1:
        assert got == omsm.msm_pippenger(g2_pts(bases), s, c=4, g2=True)


def test_msm_adversarial_inputs(ctx377):
    rng = Xoshiro(SEED_BASE + 66)
    n = 4096
    k = fr_rand(rng, n)
    bases = ctx377.g1_fixed_base_mul(fr_arr(k))
    # witness-like scalars: mostly 0/1, some r-1 and carries through every window, infinity bases in between
    s = [(0, 1, R - 1, (1 << 252) - 1, 1 << 15, (1 << 16) - 1, R - 2, 2)[i % 8] for i in range(n)]
    bases[5::97] = 0
    kk = [0 if (i % 97 == 5) else v for i, v in enumerate(k)]
    assert g1_pts(ctx377.msm_g1(fr_arr(s), bases))[0] == C.g1_mul(C.G1_GEN, dot(s, kk))
    # identical bases (DummySetup-like keys, setup.go:544-558) with random scalars
    same = np.repeat(bases[:1], n, axis=0)
    s2 = fr_rand(rng, n)
    assert g1_pts(ctx377.msm_g1(fr_arr(s2), same))[0] == C.g1_mul(C.G1_GEN, sum(s2) % R * k[0] % R)


@pytest.mark.parametrize("c", [8, 11, 13, 16])
def test_msm_every_window_size(ctx377, c):
    rng = Xoshiro(SEED_BASE + 67)
    n = 20000
    k, s = fr_rand(rng, n), fr_rand(rng, n)
    bases = ctx377.g1_fixed_base_mul(fr_arr(k))
    ctx377.set("msm_c", c)
    try:
        got = g1_pts(ctx377.msm_g1(fr_arr(s), bases))[0]
    finally:
        ctx377.set("msm_c", 0)
    assert got == C.g1_mul(C.G1_GEN, dot(s, k))


def test_msm_g1_2p18_trapdoor(ctx377):
    import torch
    n = 1 << 18
    rng = np.random.default_rng(68)
    k = rng.integers(0, 1 << 62, size=(n, 4), dtype=np.uint64)      # Montgomery images < 2^254 < ... reduce below
    k[:, 3] &= np.uint64((1 << 58) - 1)                              # < r as integers: valid Montgomery residues
    s = rng.integers(0, 1 << 62, size=(n, 4), dtype=np.uint64)
    s[:, 3] &= np.uint64((1 << 58) - 1)
    kd = torch.from_numpy(k.view(np.int64)).cuda()
    sd = torch.from_numpy(s.view(np.int64)).cuda()
    bases = torch.empty((n, 12), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    ctx377.g1_fixed_base_mul_dev(kd.data_ptr(), n, bases.data_ptr())
    got = g1_pts(ctx377.msm_g1_dev(sd.data_ptr(), bases.data_ptr(), n))[0]
    assert got == C.g1_mul(C.G1_GEN, dot(fr_ints(s), fr_ints(k)))


# ---------------------------------------------------------------- prove
def _pk_arrays(pk_py, nb_wires):
    return g16.ProvingKey(
        domain_n=pk_py.domain.n, g1_alpha=g1_arr([pk_py.g1_alpha])[0], g1_beta=g1_arr([pk_py.g1_beta])[0],
        g1_delta=g1_arr([pk_py.g1_delta])[0], g1_a=g1_arr(pk_py.g1_a), g1_b=g1_arr(pk_py.g1_b), g1_z=g1_arr(pk_py.g1_z),
        g1_k=g1_arr(pk_py.g1_k), g2_beta=g2_arr([pk_py.g2_beta])[0], g2_delta=g2_arr([pk_py.g2_delta])[0],
        g2_b=g2_arr(pk_py.g2_b), infinity_a=np.array(pk_py.infinity_a, dtype=np.uint8),
        infinity_b=np.array(pk_py.infinity_b, dtype=np.uint8), nb_wires=nb_wires, nb_public=pk_py.nb_public, curve=CURVE)


def _solve_chain(cs, rng):
    w = orcs.mimc_chain_witness(cs, rng.field(R), rng.field(R))
    w[1] = sum(cf * w[i] for cf, i in cs.constraints[-1][0]) % R
    return w


@pytest.mark.parametrize("case", ["cubic", "chain17_200"])
def test_prove_matches_oracle_bytes(case):
    rng = Xoshiro(SEED_BASE + 69)
    if case == "cubic":
        cs, w = orcs.cubic_circuit(), orcs.cubic_witness(3, 35)
    else:
        cs, _ = orcs.mimc_chain(200, rng, power=17)
        w = _solve_chain(cs, rng)
    pk_py, vk_py, _ = og.setup(cs, rng)
    r, s = rng.field(R), rng.field(R)
    parts = og.prove_parts(cs, pk_py, w)
    want = og.assemble(pk_py, parts, r, s)
    a, b, c = cs.abc(w)
    pk = _pk_arrays(pk_py, cs.nb_wires)
    try:
        sol = g16.R1CSSolution(W=fr_arr(w), A=fr_arr(a), B=fr_arr(b), C=fr_arr(c))
        points, _ = g16.prove_msms(pk, sol)
        assert g1_pts(points[:48]) == [parts["ar"], parts["bs1"], parts["krs_k"], parts["krs_hz"]]
        assert g2_pts(points[48:]) == [parts["bs2"]]
        proof = g16.Prove(None, pk, sol, g16.WithIcicleAcceleration(), g16.WithRandomness(fr_arr([r])[0], fr_arr([s])[0]))
        assert proof.raw() == og.proof_bytes(want)
        # and the GPU's proof passes the restated pairing check of verify.go (oracle.bls12_377.pairing)
        raw = proof.raw()
        got = {"ar": C.g1_from_bytes(raw[:96]), "bs": C.g2_from_bytes(raw[96:288]), "krs": C.g1_from_bytes(raw[288:384])}
        assert og.verify(got, vk_py, w[1:cs.nb_public])
        assert not og.verify(got, vk_py, [(x + 1) % R for x in w[1:cs.nb_public]])
    finally:
        pk.free_device()


def test_prove_2p13_trapdoor(ctx377):
    """pow17 chain of 2^13 constraints: key = [setup scalars]G from the fixed-base kernel, expected MSM points on the
    exponent side, h from the oracle."""
    rng = Xoshiro(SEED_BASE + 70)
    nc = 1 << 13
    cs, _ = orcs.mimc_chain(nc, rng, power=17)
    w = _solve_chain(cs, rng)
    d = ontt.Domain(nc)
    tw = og.ToxicWaste(rng)
    A, B, Cc = og.setup_abc(cs, d, tw)
    nw, npub = cs.nb_wires, cs.nb_public
    inf_a = [v == 0 for v in A]
    inf_b = [v == 0 for v in B]
    sA = [v for v in A if v]
    sB = [v for v in B if v]
    sK = [(tw.beta * A[i] + tw.alpha * B[i] + Cc[i]) % R * tw.delta_inv % R for i in range(npub, nw)]
    zt = (pow(tw.t, nc, R) - 1) * tw.delta_inv % R
    Z = [zt * pow(tw.t, i, R) % R for i in range(nc)]
    sZ = ontt.bit_reverse_list(Z)[: nc - 1]
    g1 = ctx377.g1_fixed_base_mul(fr_arr([tw.alpha, tw.beta, tw.delta] + sA + sB + sK + sZ))
    g2 = ctx377.g2_fixed_base_mul(fr_arr([tw.beta, tw.delta] + sB))
    o = 3
    pa = g1[o:o + len(sA)]; o += len(sA)
    pb = g1[o:o + len(sB)]; o += len(sB)
    pkk = g1[o:o + len(sK)]; o += len(sK)
    pz = g1[o:]
    pk = g16.ProvingKey(domain_n=nc, g1_alpha=g1[0], g1_beta=g1[1], g1_delta=g1[2], g1_a=pa, g1_b=pb, g1_z=pz, g1_k=pkk,
                        g2_beta=g2[0], g2_delta=g2[1], g2_b=g2[2:], infinity_a=np.array(inf_a, dtype=np.uint8),
                        infinity_b=np.array(inf_b, dtype=np.uint8), nb_wires=nw, nb_public=npub, curve=CURVE)
    a, b, c = cs.abc(w)
    h = ontt.compute_h(a, b, c, d)
    wa = [w[i] for i in range(nw) if not inf_a[i]]
    wb = [w[i] for i in range(nw) if not inf_b[i]]
    try:
        sol = g16.R1CSSolution(W=fr_arr(w), A=fr_arr(a), B=fr_arr(b), C=fr_arr(c))
        points, st = g16.prove_msms(pk, sol)
        want_g1 = [C.g1_mul(C.G1_GEN, e) for e in (dot(wa, sA), dot(wb, sB), dot(w[npub:], sK), dot(h[: nc - 1], sZ))]
        assert g1_pts(points[:48]) == want_g1
        assert g2_pts(points[48:]) == [C.g2_mul(C.G2_GEN, dot(wb, sB))]
        # 2-way point-range shards add up to the same points
        blocks = []
        for si in range(2):
            pk.free_device()
            pk.setup_device_pointers(g16.NewProverConfig(g16.WithDevice(0, si, 2)))
            blocks.append(g16.prove_msms(pk, sol)[0])
        assert np.array_equal(g16.combine_shards(blocks, curve=CURVE), points)
    finally:
        pk.free_device()


# ---------------------------------------------------------------- BSB22 commitment keys, resident bases (8(f) N3)
@pytest.mark.parametrize("n", [0, 1, 40, 900])
def test_commit_matches_oracle(ctx377, n):
    """g16_bls12_377_commit: commitment = sum v_j Basis_j, pok = sum v_j BasisExpSigma_j (12-limb Fp build)."""
    rng = Xoshiro(SEED_BASE + 75 + n)
    logs = fr_rand(rng, n)
    sigma = rng.field(R) or 1
    vals = fr_rand(rng, n)
    if n > 2:
        vals[0] = 0
        logs[1] = 0                                  # an infinity base is legal
    empty = np.zeros((0, 12), dtype=np.uint64)      # bases from the fixed-base kernel (itself checked against the oracle)
    basis = ctx377.g1_fixed_base_mul(fr_arr(logs)) if n else empty
    basis_sigma = ctx377.g1_fixed_base_mul(fr_arr([k * sigma % R for k in logs])) if n else empty
    h = ctx377.commitment_key_upload(basis, basis_sigma)
    cm, pok = ctx377.commit(h, fr_arr(vals) if n else np.zeros((0, 4), np.uint64))
    e = dot(vals, logs)
    assert g1_pts(cm)[0] == (C.g1_mul(C.G1_GEN, e) if e else None)
    assert g1_pts(pok)[0] == (C.g1_mul(C.G1_GEN, e * sigma % R) if e else None)
    if n:
        with pytest.raises(lib.G16Error, match="as many values as basis elements"):
            ctx377.commit(h, fr_arr(vals[:-1]) if n > 1 else np.zeros((0, 4), np.uint64))
    ctx377.commitment_key_free(h)


def test_msm_resident_host_scalars_g1_g2(ctx377):
    rng = Xoshiro(SEED_BASE + 76)
    logs, vals = fr_rand(rng, 600), fr_rand(rng, 600)
    e = dot(vals, logs)
    h1 = ctx377.bases_upload(ctx377.g1_fixed_base_mul(fr_arr(logs)))
    assert g1_pts(ctx377.msm_resident(h1, fr_arr(vals)))[0] == C.g1_mul(C.G1_GEN, e)
    ctx377.bases_free(h1)
    h2 = ctx377.bases_upload(ctx377.g2_fixed_base_mul(fr_arr(logs)), g2=True)
    assert g2_pts(ctx377.msm_resident(h2, fr_arr(vals), g2=True))[0] == C.g2_mul(C.G2_GEN, e)
    ctx377.bases_free(h2)
