"""Proof wire format (backend/groth16/bn254/marshal.go:30-92): the product's WriteTo / WriteRawTo /
ReadFrom against the oracle's independent writer, on proofs with and without BSB22 commitments.  CPU only."""
import numpy as np
import pytest

import gnark_icicle_b200 as g16
from gnark_icicle_b200 import marshal
from oracle import groth16 as G, r1cs as RC, curve as C, marshal as omarshal
from oracle import cpu as ocpu
from oracle.prng import Xoshiro


def _proof(ncm, seed):
    rng = Xoshiro(seed)
    if ncm:
        cs = RC.committed_chain(30, rng, ncm)
        pk, vk, _ = G.setup(cs, rng)
        asg = RC.committed_chain_assignment(cs, pk, 3, 4, 5, G.solve)
        proof, _ = G.prove_full(cs, pk, asg, rng.field(C.R), rng.field(C.R))
    else:
        cs = RC.cubic_circuit()
        pk, vk, _ = G.setup(cs, rng)
        proof = G.prove(cs, pk, RC.cubic_witness(3, 35), rng.field(C.R), rng.field(C.R))
        proof["commitments"], proof["commitment_pok"] = [], None
    g1 = lambda p: ocpu.g1_arr([p])[0]
    mirror = g16.Proof(Ar=g1(proof["ar"]), Bs=ocpu.g2_arr([proof["bs"]])[0], Krs=g1(proof["krs"]),
                       Commitments=[g1(c) for c in proof["commitments"]],
                       CommitmentPok=g1(proof["commitment_pok"]))
    return proof, mirror


@pytest.mark.parametrize("ncm", [0, 1, 3])
@pytest.mark.parametrize("raw", [False, True])
def test_proof_write_to_matches_oracle_and_round_trips(ncm, raw):
    for seed in (11, 12, 13):       # both y signs turn up across seeds
        want, mirror = _proof(ncm, seed)
        data = marshal.proof_write_to(mirror, raw=raw)
        assert data == omarshal.proof_bytes(want, raw=raw)
        assert len(data) == (256 if raw else 128) + 4 + (ncm + 1) * (64 if raw else 32)
        back, n = marshal.proof_read_from(data)
        assert n == len(data)
        assert back.raw() == mirror.raw() if ncm else back.raw()[:256] == mirror.raw()[:256]
        assert len(back.Commitments) == ncm


def test_decode_rejects_points_off_the_curve():
    _, mirror = _proof(0, 11)
    data = bytearray(marshal.proof_write_to(mirror, raw=True))
    data[63] ^= 1                       # Ar.Y
    with pytest.raises(ValueError, match="not on the curve"):
        marshal.proof_read_from(bytes(data))


# ---------------------------------------------------------------- VerifyingKey / ProvingKey (marshal.go:99-236, :242-385)
def _keys(ncm, seed):
    from helpers import pk_from_python
    from gnark_icicle_b200 import bsb22
    rng = Xoshiro(seed)
    cs = RC.committed_chain(30, rng, ncm) if ncm else RC.cubic_circuit()
    pk, vk, _ = G.setup(cs, rng)
    mpk = pk_from_python(pk, cs.nb_wires)
    mpk.commitment_keys = [bsb22.CommitmentKey(ocpu.g1_arr(k.basis) if k.basis else np.zeros((0, 8), dtype=np.uint64),
                                                ocpu.g1_arr(k.basis_exp_sigma) if k.basis else np.zeros((0, 8), dtype=np.uint64))
                           for k in pk.commitment_keys]
    g1 = lambda p: ocpu.g1_arr([p])[0]
    g2 = lambda p: ocpu.g2_arr([p])[0]
    mvk = marshal.VerifyingKey(g1(vk.g1_alpha), g1(pk.g1_beta), g2(vk.g2_beta), g2(vk.g2_gamma), g1(pk.g1_delta), g2(vk.g2_delta),
                               ocpu.g1_arr(vk.g1_k), [list(r) for r in vk.public_and_commitment_committed],
                               [marshal.PedersenVerifyingKey(g2(k.g), g2(k.g_sigma_neg)) for k in vk.commitment_keys])
    vk.g1_beta, vk.g1_delta = pk.g1_beta, pk.g1_delta
    return pk, vk, mpk, mvk


@pytest.mark.parametrize("ncm", [0, 2])
@pytest.mark.parametrize("raw", [False, True])
def test_key_wire_formats_match_oracle_and_round_trip(ncm, raw):
    pk, vk, mpk, mvk = _keys(ncm, 41 + ncm)
    data = marshal.vk_write_to(mvk, raw=raw)
    assert data == omarshal.vk_bytes(vk, raw=raw)
    back, n = marshal.vk_read_from(data)
    assert n == len(data) and marshal.vk_write_to(back, raw=raw) == data
    assert back.public_and_commitment_committed == mvk.public_and_commitment_committed
    data = marshal.pk_write_to(mpk, raw=raw)
    assert data == omarshal.pk_bytes(pk, raw=raw)
    back, n = marshal.pk_read_from(data, nb_public=mpk.nb_public)
    assert n == len(data)
    for f in ("g1_alpha", "g1_beta", "g1_delta", "g1_a", "g1_b", "g1_z", "g1_k", "g2_beta", "g2_delta", "g2_b", "infinity_a", "infinity_b"):
        assert np.array_equal(np.asarray(getattr(back, f)).reshape(-1), np.asarray(getattr(mpk, f)).reshape(-1)), f
    assert len(back.commitment_keys) == ncm
    for a, b in zip(back.commitment_keys, mpk.commitment_keys):
        assert np.array_equal(a.basis, b.basis) and np.array_equal(a.basis_exp_sigma, b.basis_exp_sigma)
    # the other encoding of the same key decodes to the same key (ReadFrom accepts both)
    other, _ = marshal.pk_read_from(marshal.pk_write_to(mpk, raw=not raw), nb_public=mpk.nb_public)
    assert marshal.pk_write_to(other, raw=raw) == data
