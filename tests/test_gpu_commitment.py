"""BSB22 commitments on the device: g16_commit against the oracle's Pedersen MSMs, and Prove for
circuits that commit (the shape of backend/groth16/bn254/commitment_test.go) -- proof bytes, commitments and the
folded proof of knowledge bit-exact against the oracle prover, and accepted by the restated Verify."""
import numpy as np
import pytest

import gnark_icicle_b200 as g16
from gnark_icicle_b200 import bsb22, lib
from oracle import commitment as CM, groth16 as G, r1cs as RC, curve as C
from oracle import cpu as ocpu
from oracle.prng import Xoshiro

from helpers import fr_rand, g1_bases, pk_from_python, trapdoor_g1, assert_limbs_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = lib.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("n", [0, 1, 2, 7, 300, 5000, 40000])
def test_commit_matches_oracle(ctx, n):
    """commitment = sum v_j Basis_j, pok = sum v_j BasisExpSigma_j (pedersen Commit / ProveKnowledge); 40000 is
    above the threshold where the key gets its table of precomputed multiples."""
    if n == 0:
        h = ctx.commitment_key_upload(np.zeros((0, 8), np.uint64), np.zeros((0, 8), np.uint64))
        cm, pok = ctx.commit(h, np.zeros((0, 4), np.uint64))
        assert not cm.any() and not pok.any()
        ctx.commitment_key_free(h)
        return
    logs, basis = g1_bases(300 + n, n)
    sigma = fr_rand(77, 1)
    logs_sigma = ocpu.fr_mul_vec(logs, np.repeat(sigma, n, axis=0))
    basis_sigma = ocpu.g1_mul_gen_batch(logs_sigma)
    vals = fr_rand(500 + n, n)
    vals[0] = 0                      # a zero value and, below, an infinity base are legal
    if n > 2:
        basis[1] = 0
        basis_sigma[1] = 0
        logs[1] = 0
        logs_sigma[1] = 0
    h = ctx.commitment_key_upload(basis, basis_sigma)
    cm, pok = ctx.commit(h, vals)
    assert_limbs_equal(cm, trapdoor_g1(vals, logs), "commitment")
    assert_limbs_equal(pok, trapdoor_g1(vals, logs_sigma), "pok")
    if n <= 5000:
        assert_limbs_equal(cm, ocpu.msm_g1(basis, vals), "commitment vs CPU Pippenger")
    cm2, none = ctx.commit(h, vals, want_pok=False)
    assert none is None and np.array_equal(cm, cm2)
    with pytest.raises(lib.G16Error, match="as many values as basis elements"):
        ctx.commit(h, vals[:-1])
    ctx.commitment_key_free(h)
    with pytest.raises(lib.G16Error, match="unknown commitment-key handle"):
        ctx.commit(h, vals)


def test_msm_resident_host_scalars(ctx):
    logs, basis = g1_bases(901, 3000)
    vals = fr_rand(902, 3000)
    h = ctx.bases_upload(basis)
    assert_limbs_equal(ctx.msm_resident(h, vals), trapdoor_g1(vals, logs), "g16_msm_resident")
    ctx.bases_free(h)


class OracleR1CS:
    """what Prove needs from cs.R1CS for a circuit with commitments; the solver is the oracle's (host code outside
    the accelerated path), with the bsb22 hint handed in by Prove like solver.OverrideHint (prove.go:81-83)."""

    def __init__(self, cs, pk_py):
        self.cs, self.pk_py = cs, pk_py
        self.nb_public = cs.nb_public
        self.commitment_info = [bsb22.Groth16Commitment(c.public_and_commitment_committed, c.private_committed,
                                                        c.commitment_index, c.nb_public_committed)
                                for c in cs.commitment_info]

    def solve(self, assignment, bsb22_hint):
        w, _, _ = G.solve(self.cs, self.pk_py, assignment, hint_fn=bsb22_hint)
        a, b, c = self.cs.abc(w)
        return g16.R1CSSolution(W=ocpu.fr_arr(w), A=ocpu.fr_arr(a), B=ocpu.fr_arr(b), C=ocpu.fr_arr(c))


def _pk(pk_py, cs):
    pk = pk_from_python(pk_py, cs.nb_wires)
    g1 = lambda pts: ocpu.g1_arr(pts) if len(pts) else np.zeros((0, 8), dtype=np.uint64)
    pk.commitment_keys = [bsb22.CommitmentKey(g1(k.basis), g1(k.basis_exp_sigma)) for k in pk_py.commitment_keys]
    return pk


@pytest.mark.parametrize("nc,ncm", [(20, 1), (40, 2), (50, 3), (700, 2)])
def test_prove_with_commitments_matches_oracle(nc, ncm):
    rng = Xoshiro(100 + nc)
    cs = RC.committed_chain(nc, rng, ncm, nb_private_committed=5 if nc < 100 else 200)
    pk_py, vk, scal = G.setup(cs, rng)
    asg = RC.committed_chain_assignment(cs, pk_py, 5, 6, 7, G.solve)
    r, s = rng.field(C.R), rng.field(C.R)
    want, w = G.prove_full(cs, pk_py, asg, r, s)
    pk = _pk(pk_py, cs)
    try:
        proof = g16.Prove(OracleR1CS(cs, pk_py), pk, asg, g16.WithIcicleAcceleration(),
                          g16.WithRandomness(ocpu.fr_arr([r])[0], ocpu.fr_arr([s])[0]))
        assert len(proof.Commitments) == ncm
        for i in range(ncm):
            assert_limbs_equal(proof.Commitments[i], ocpu.g1_arr([want["commitments"][i]])[0], f"commitment {i}")
        assert_limbs_equal(proof.CommitmentPok, ocpu.g1_arr([want["commitment_pok"]])[0], "folded pok")
        assert proof.raw()[:256] == G.proof_bytes(want), "Ar || Bs || Krs differ from the oracle prover"
        got = dict(ar=ocpu.g1_pt(proof.Ar), bs=ocpu.g2_pt(proof.Bs), krs=ocpu.g1_pt(proof.Krs),
                   commitments=[ocpu.g1_pt(c) for c in proof.Commitments], commitment_pok=ocpu.g1_pt(proof.CommitmentPok))
        if nc <= 40:
            assert G.verify(got, vk, asg[1:3])
        # a second proof against the resident keys
        proof2 = g16.Prove(OracleR1CS(cs, pk_py), pk, asg, g16.WithIcicleAcceleration(),
                           g16.WithRandomness(ocpu.fr_arr([r])[0], ocpu.fr_arr([s])[0]))
        assert proof2.raw() == proof.raw()
    finally:
        pk.free_device()


def test_prove_with_commitments_error_paths():
    rng = Xoshiro(55)
    cs = RC.committed_chain(20, rng, 1)
    pk_py, vk, scal = G.setup(cs, rng)
    asg = RC.committed_chain_assignment(cs, pk_py, 5, 6, 7, G.solve)
    pk = _pk(pk_py, cs)
    r1 = OracleR1CS(cs, pk_py)
    try:
        with pytest.raises(lib.G16Error, match="solved inside Prove"):
            g16.Prove(None, pk, g16.R1CSSolution(None, None, None, None), g16.WithIcicleAcceleration())
        pk.commitment_keys = []
        with pytest.raises(lib.G16Error, match="commitments, the proving key"):
            g16.Prove(r1, pk, asg, g16.WithIcicleAcceleration())
    finally:
        pk.free_device()
