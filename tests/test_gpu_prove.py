"""GPU parity of the whole prove path through the plugin surface (gnark_icicle_b200.Prove) vs the oracle."""
import numpy as np
import pytest

from helpers import ChainCase, assert_limbs_equal, fr_from_ints, pk_from_python, R
from oracle import cpu as ocpu
from oracle import groth16 as og
from oracle import r1cs as orcs
from oracle.prng import Xoshiro, SEED_BASE
import gnark_icicle_b200 as g16

pytestmark = pytest.mark.gpu


def _py_case(cs, w, seed):
    rng = Xoshiro(SEED_BASE + seed)
    pk_py, vk, scal = og.setup(cs, rng)
    r, s = rng.field(R), rng.field(R)
    parts = og.prove_parts(cs, pk_py, w)
    proof_py = og.assemble(pk_py, parts, r, s)
    return pk_py, vk, parts, proof_py, r, s


def test_cubic_proof_is_byte_exact_and_verifies():
    """BASELINE config 1: examples/cubic (X=3, Y=35), fixed toxic waste and r,s."""
    cs = orcs.cubic_circuit()
    w = orcs.cubic_witness(3, 35)
    pk_py, vk, parts, proof_py, r, s = _py_case(cs, w, 1)
    pk = pk_from_python(pk_py, cs.nb_wires)
    a, b, c = cs.abc(w)
    sol = g16.R1CSSolution(W=fr_from_ints(w), A=fr_from_ints(a), B=fr_from_ints(b), C=fr_from_ints(c))
    proof = g16.Prove(None, pk, sol, g16.WithIcicleAcceleration(), g16.WithRandomness(fr_from_ints([r])[0], fr_from_ints([s])[0]))
    assert proof.raw() == og.proof_bytes(proof_py)
    got = dict(ar=ocpu.g1_pt(proof.Ar), bs=ocpu.g2_pt(proof.Bs), krs=ocpu.g1_pt(proof.Krs))
    assert og.verify(got, vk, w[1:cs.nb_public])
    assert not og.verify(got, vk, [36])
    pk.free_device()


def test_cubic_raw_msm_points_and_stats():
    cs = orcs.cubic_circuit()
    w = orcs.cubic_witness(3, 35)
    pk_py, vk, parts, proof_py, r, s = _py_case(cs, w, 2)
    pk = pk_from_python(pk_py, cs.nb_wires)
    a, b, c = cs.abc(w)
    sol = g16.R1CSSolution(W=fr_from_ints(w), A=fr_from_ints(a), B=fr_from_ints(b), C=fr_from_ints(c))
    pts, stats = g16.prove_msms(pk, sol)
    for i, name in enumerate(["ar", "bs1", "krs_k", "krs_hz"]):
        assert ocpu.g1_pt(pts[8 * i:8 * i + 8]) == parts[name], name
    assert ocpu.g2_pt(pts[32:48]) == parts["bs2"]
    assert stats["kernel_launches"] > 0 and stats["h2d_bytes"] == 32 * (5 + 3 * 3)
    pk.free_device()


def test_without_acceleration_option_is_an_error():
    with pytest.raises(g16.lib.G16Error):
        g16.Prove(None, None, None)


def test_mimc_python_oracle_proof():
    """small MiMC-shaped chain end to end against the big-int oracle, incl. pairing verification."""
    rng = Xoshiro(SEED_BASE + 31)
    cs, _ = orcs.mimc_chain(40, rng)
    w = orcs.mimc_chain_witness(cs, 123456789, 987654321)
    # the chain's last constraint binds the public output
    a, b, c = None, None, None
    L, Rr, O = cs.constraints[-1]
    w[1] = sum(cf * w[i] for cf, i in L) % R
    pk_py, vk, parts, proof_py, r, s = _py_case(cs, w, 32)
    pk = pk_from_python(pk_py, cs.nb_wires)
    a, b, c = cs.abc(w)
    sol = g16.R1CSSolution(W=fr_from_ints(w), A=fr_from_ints(a), B=fr_from_ints(b), C=fr_from_ints(c))
    proof = g16.Prove(None, pk, sol, g16.WithIcicleAcceleration(), g16.WithRandomness(fr_from_ints([r])[0], fr_from_ints([s])[0]))
    assert proof.raw() == og.proof_bytes(proof_py)
    assert og.verify(dict(ar=ocpu.g1_pt(proof.Ar), bs=ocpu.g2_pt(proof.Bs), krs=ocpu.g1_pt(proof.Krs)), vk, [w[1]])
    pk.free_device()


@pytest.mark.parametrize("nc", [1000, 1 << 14])
def test_chain_prove_matches_c_oracle(nc):
    """config-4 shape at sizes the CPU oracle finishes in seconds: h and all five MSM points, two ways."""
    case = ChainCase(nc, seed=nc % 97)
    sol = g16.R1CSSolution(W=case.W, A=case.a, B=case.b, C=case.c)
    pts, stats = g16.prove_msms(case.pk, sol)
    h = ocpu.compute_h(case.a, case.b, case.c, case.n)
    pk = case.pk
    wa, wb, wk = case.W[~case.inf_a], case.W[~case.inf_b], case.W[2:]
    want = np.concatenate([ocpu.msm_g1(pk.g1_a, wa), ocpu.msm_g1(pk.g1_b, wb), ocpu.msm_g1(pk.g1_k, wk),
                           ocpu.msm_g1(pk.g1_z, h[: case.n - 1]), ocpu.msm_g2(pk.g2_b, wb)])
    assert_limbs_equal(pts, want, "prove points vs CPU Pippenger")
    assert_limbs_equal(pts, case.expected_points(h), "prove points vs trapdoor identity")
    pk.free_device()


def test_sharded_prove_combines_to_the_same_points():
    """point-range sharding: shard_count 3 run back to back on one GPU == unsharded."""
    case = ChainCase(3000, seed=5)
    sol = g16.R1CSSolution(W=case.W, A=case.a, B=case.b, C=case.c)
    whole, _ = g16.prove_msms(case.pk, sol)
    ctx0, h0 = case.pk.device
    assert ctx0.pk_wire_range(h0) == (0, case.nw)        # unsharded: the A map starts at wire 0, K ends at the last wire
    case.pk.free_device()
    blocks, ranges = [], []
    for i in range(3):
        cfg = g16.NewProverConfig(g16.WithDevice(0, i, 3))
        # the host-buffer entry point uploads only the wires this shard reads: poison everything else
        blk, st = g16.prove_msms(case.pk, sol, cfg)
        ctx_i, h_i = case.pk.device
        lo, hi = ctx_i.pk_wire_range(h_i)
        ranges.append((lo, hi))
        assert st["h2d_bytes"] == (hi - lo) * 32 + 3 * len(case.a) * 32
        Wp = case.W.copy()
        Wp[:lo] = 0xDEADBEEF
        Wp[hi:] = 0xDEADBEEF
        blk2, _ = g16.prove_msms(case.pk, g16.R1CSSolution(W=Wp, A=case.a, B=case.b, C=case.c), cfg)
        assert np.array_equal(blk, blk2), "wires outside the shard's range must not be read"
        blocks.append(blk)
        case.pk.free_device()
    assert_limbs_equal(g16.combine_shards(blocks), whole, "sharded vs whole")
    assert ranges[0][0] == 0 and ranges[-1][1] == case.nw and all(hi - lo < case.nw * 0.4 for lo, hi in ranges)


def test_full_size_prove_2p20_trapdoor(ctx):
    """2^20 constraints (one step below config 4 to keep the suite in minutes): key materialised by the GPU
    fixed-base kernel from the oracle's trapdoor scalars; all five points bit-exact via the trapdoor identity,
    final proof satisfies the exponent-side Groth16 equation."""
    nc = 1 << 20
    case = ChainCase(nc, seed=7, g1_mul=ctx.g1_fixed_base_mul, g2_mul=ctx.g2_fixed_base_mul)
    sol = g16.R1CSSolution(W=case.W, A=case.a, B=case.b, C=case.c)
    pts, stats = g16.prove_msms(case.pk, sol)
    h = ocpu.compute_h(case.a, case.b, case.c, case.n)
    assert_limbs_equal(pts, case.expected_points(h), "2^20 prove points vs trapdoor identity")
    case.pk.free_device()


# ---------------------------------------------------------------- precomputed-multiples mode (shared bucket set)
@pytest.mark.parametrize("kind", ["random", "bits", "equal", "sharded"])
def test_precomputed_tables_give_the_same_points_as_the_window_path(ctx, kind):
    """g16_pk_upload keeps 2^(c j) multiples of every base (n >= 2^14) and the prove MSMs then use ONE bucket set with
    a two-level sort; "msm_precompute" = 0 after the upload runs the per-window path on the same key.  Both must give
    identical points (and both are checked against the trapdoor in the other tests)."""
    nc = 1 << 15
    case = ChainCase(nc, seed=41, g1_mul=ctx.g1_fixed_base_mul, g2_mul=ctx.g2_fixed_base_mul)
    W = case.W.copy()
    if kind == "bits":          # witness-like: most wires are 0 / 1, a few r-1
        W[4::2] = 0
        W[5::4] = ocpu.fr_arr([1])[0]
        W[7::64] = ocpu.fr_arr([R - 1])[0]
    elif kind == "equal":       # every wire the same value: all entries of a window land in one bucket
        W[:] = W[9]
    sol = g16.R1CSSolution(W=W, A=case.a, B=case.b, C=case.c)
    pk = case.pk
    shards = 3 if kind == "sharded" else 1
    try:
        got, want = [], []
        for si in range(shards):
            pk.free_device()
            pk.setup_device_pointers(g16.NewProverConfig(g16.WithDevice(0, si, shards)))
            c = pk.device[0]
            p1, st1 = g16.prove_msms(pk, sol)
            c.set("msm_precompute", 0)
            p0, st0 = g16.prove_msms(pk, sol)
            if shards == 1:
                assert st1["msm_windows"][0] < st0["msm_windows"][0], (st1["msm_windows"], st0["msm_windows"])
            got.append(p1)
            want.append(p0)
        assert np.array_equal(g16.combine_shards(got), g16.combine_shards(want))
        if kind == "random":
            h = ocpu.compute_h(case.a, case.b, case.c, case.n)
            assert_limbs_equal(got[0], case.expected_points(h), "precomputed mode vs trapdoor")
    finally:
        pk.free_device()


# ---------------------------------------------------------------- G1.K lined up with G1.A (shared sort)
@pytest.mark.parametrize("shape", ["tail", "tail_pad4", "hole", "k_longer", "k_index"])
def test_krs_msm_reusing_the_ar_sort_matches_oracle(ctx, shape):
    """When the wires of G1.K are a tail of the wires of G1.A, g16_pk_upload pads G1.K with points at infinity so that the
    two arrays line up and the krs MSM reuses the ar MSM's digits and bucket sort (csrc/api.cu pk_upload_ranges).  All
    five points against the oracle's bucket-method MSMs for keys where that applies (pad 1 as in every gnark key with
    one public input; pad 4) and for keys where it must not (a wire missing from A in the middle; K longer than A; an
    explicit k_index map)."""
    nc = (1 << 14) + 37
    case = ChainCase(nc, seed=73)
    pk = case.pk
    nw = case.nw
    inf_a, inf_b = case.inf_a.copy(), case.inf_b.copy()
    g1_a, g1_k, k_index = pk.g1_a, pk.g1_k, None
    wk = case.W[2:]
    if shape == "tail_pad4":            # five public wires: K starts at wire 5, A still lacks only wire 1
        g1_k, wk = pk.g1_k[3:], case.W[5:]
    elif shape == "hole":               # wire 100 has a zero A polynomial: same lengths, maps differ from there on
        pos = int(np.count_nonzero(~inf_a[:100]))
        inf_a[100] = True
        g1_a = np.delete(pk.g1_a, pos, axis=0)
    elif shape == "k_longer":           # wires 5..9 missing from A: K (nw - 2 points) is longer than A (nw - 6)
        pos = int(np.count_nonzero(~inf_a[:5]))
        inf_a[5:10] = True
        g1_a = np.delete(pk.g1_a, np.arange(pos, pos + 5), axis=0)
    elif shape == "k_index":            # an explicit map (what a key with commitments passes): every other private wire
        k_index = np.arange(2, nw, 2, dtype=np.uint32)
        g1_k, wk = pk.g1_k[::2], case.W[2::2]
        assert len(g1_k) == len(k_index)
    handle = ctx.pk_upload(pk.domain_n, g1_a, pk.g1_b, g1_k, pk.g1_z, pk.g2_b, inf_a.astype(np.uint8), inf_b.astype(np.uint8), nw,
                           k_index=k_index)
    try:
        pts, st = ctx.prove(handle, case.W, case.a, case.b, case.c)
        pts2, _ = ctx.prove(handle, case.W, case.a, case.b, case.c)
    finally:
        ctx.pk_free(handle)
    h = ocpu.compute_h(case.a, case.b, case.c, case.n)
    wa, wb = case.W[~inf_a], case.W[~inf_b]
    want = np.concatenate([ocpu.msm_g1(g1_a, wa), ocpu.msm_g1(pk.g1_b, wb), ocpu.msm_g1(g1_k, wk),
                           ocpu.msm_g1(pk.g1_z, h[: case.n - 1]), ocpu.msm_g2(pk.g2_b, wb)])
    assert_limbs_equal(pts, want, f"{shape}: prove points vs CPU Pippenger")
    assert_limbs_equal(pts2, pts, f"{shape}: second prove on the same key")
    assert list(st["msm_n"][:3]) == [len(g1_a), len(pk.g1_b), len(g1_k)]     # the padding does not show in the stats
