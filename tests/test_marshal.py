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
