"""BSB22 commitments: the oracle restatement against the vectors that exist for it, and the
host-side pieces of the product (hash to field, Marshal, K-wire filtering, folding) against the oracle.  CPU only."""
import numpy as np
import pytest

from oracle import commitment as CM, groth16 as G, r1cs as RC, curve as C
from oracle import cpu as ocpu
from oracle.prng import Xoshiro

from gnark_icicle_b200 import bsb22, lib

# RFC 9380 appendix K.1 (expand_message_xmd, SHA-256, DST below): the vectors gnark-crypto's own hash_to_field tests use
RFC_DST = b"QUUX-V01-CS02-with-expander-SHA256-128"
RFC_VECTORS = [
    (b"", 0x20, "68a985b87eb6b46952128911f2a4412bbc302a9d759667f87f7a21d803f07235"),
    (b"abc", 0x20, "d8ccab23b5985ccea865c6c97b6e5b8350e794e603b4b97902f53a8a0d605615"),
    (b"", 0x80, "af84c27ccfd45d41914fdff5df25293e221afc53d8ad2ac06d5e3e29485dadbee0d121587713a3e0dd4d5e69e93eb7cd4f5df4cd1"
                "03e188cf60cb02edc3edf18eda8576c412b18ffb658e3dd6ec849469b979d444cf7b26911a08e63cf31f9dcc541708d3491184472"
                "c2c29bb749d4286b004ceb5ee6b9a7fa5b646c993f0ced"),
]


@pytest.mark.parametrize("impl", [CM.expand_msg_xmd, bsb22.expand_msg_xmd], ids=["oracle", "product"])
def test_expand_msg_xmd_rfc9380_vectors(impl):
    for msg, n, want in RFC_VECTORS:
        assert impl(msg, RFC_DST, n).hex() == want


def test_product_hash_and_marshal_match_oracle():
    rng = Xoshiro(3)
    for n in (0, 1, 31, 32, 33, 200):
        msg = bytes(rng.next() & 0xFF for _ in range(n))
        assert bsb22.fr_hash(msg, b"G16-BSB22", 2) == CM.fr_hash(msg, b"G16-BSB22", 2)
    h1, h2 = bsb22.HashToField(), CM.HashToField()
    h1.write(b"abc"); h1.write(b"def"); h2.write(b"abcdef")
    assert h1.sum() == h2.sum() and h1.size() == h2.size() == 32
    pts = [C.g1_mul(C.G1_GEN, rng.field(C.R)) for _ in range(3)] + [None]
    for pt in pts:
        assert bsb22.g1_marshal(ocpu.g1_arr([pt])[0]) == CM.g1_marshal(pt)
    assert CM.g1_marshal(None)[0] == 0x40 and not any(CM.g1_marshal(None)[1:])
    vals = [rng.field(C.R) for _ in range(5)]
    assert np.array_equal(bsb22.fr_vec_to_limbs(vals), ocpu.fr_arr(vals))


def _case(nc, ncm, seed):
    rng = Xoshiro(seed)
    cs = RC.committed_chain(nc, rng, ncm)
    pk, vk, scal = G.setup(cs, rng)
    asg = RC.committed_chain_assignment(cs, pk, 11 + seed, 22, 33, G.solve)
    return rng, cs, pk, vk, scal, asg


@pytest.mark.parametrize("nc,ncm", [(20, 1), (40, 2), (50, 3)])
def test_oracle_prove_with_commitments_verifies(nc, ncm):
    """commitment_test.go:33-149 shape: a proof of a circuit that commits must pass Verify; the commitment, its
    proof of knowledge and the public inputs are all bound."""
    rng, cs, pk, vk, scal, asg = _case(nc, ncm, 7)
    assert len(cs.commitment_info) == ncm and len(pk.commitment_keys) == ncm
    assert len(vk.g1_k) == cs.nb_public + ncm
    assert len(pk.g1_k) == cs.nb_wires - cs.nb_public - ncm - sum(len(c.private_committed) for c in cs.commitment_info)
    proof, w = G.prove_full(cs, pk, asg, rng.field(C.R), rng.field(C.R))
    assert G.verify(proof, vk, asg[1:3])
    bad = dict(proof); bad["commitment_pok"] = C.g1_add(proof["commitment_pok"], C.G1_GEN)
    assert not G.verify(bad, vk, asg[1:3])
    bad = dict(proof); bad["commitments"] = [C.g1_add(proof["commitments"][0], C.G1_GEN)] + proof["commitments"][1:]
    assert not G.verify(bad, vk, asg[1:3])
    assert not G.verify(proof, vk, [asg[1], (asg[2] + 1) % C.R])
    # exponent side: commitment_i = [sum v_j ckK_j] G, pok_i = sigma_i * commitment_i
    for i, c in enumerate(cs.commitment_info):
        e = sum(w[j] * k for j, k in zip(c.private_committed, scal["CK"][i])) % C.R
        assert proof["commitments"][i] == C.g1_mul(C.G1_GEN, e)
    ch = CM.fold_challenge([w[c.commitment_index] for c in cs.commitment_info])
    assert proof["commitment_pok"] == CM.fold([C.g1_mul(cm, s) for cm, s in zip(proof["commitments"], scal["sigma"])], ch)


def test_pedersen_setup_literal_equals_trapdoor():
    rng, cs, pk, vk, scal, asg = _case(20, 1, 9)
    lit_pk, lit_vk = CM.pedersen_setup(pk.commitment_keys[0].basis, vk.commitment_keys[0].g, scal["sigma"][0])
    assert lit_pk.basis_exp_sigma == pk.commitment_keys[0].basis_exp_sigma
    assert lit_vk.g_sigma_neg == vk.commitment_keys[0].g_sigma_neg


def test_public_and_commitment_committed_translation():
    """constraint/commitment.go:53-74: commitment wires are renumbered to nbPublic + their rank."""
    info = [CM.Groth16Commitment([2], [3, 4], 17, 1), CM.Groth16Commitment([1, 2, 17], [20], 30, 2),
            CM.Groth16Commitment([17, 30], [], 43, 0)]
    assert CM.get_public_and_commitment_committed(info, [17, 30, 43], 3) == [[2], [1, 2, 3], [3, 4]]


def test_product_k_wire_indices_and_fold_match_oracle():
    rng, cs, pk, vk, scal, asg = _case(40, 2, 5)
    info = [bsb22.Groth16Commitment(c.public_and_commitment_committed, c.private_committed, c.commitment_index,
                                    c.nb_public_committed) for c in cs.commitment_info]
    assert list(bsb22.k_wire_indices(cs.nb_public, cs.nb_wires, info)) == G.k_wire_indices(cs)
    # filterHeap KAT shape (utils_test.go:17-38): removal lists may be unsorted and hold duplicates
    assert G.filter_heap([1, 2, 3, 4, 5], 10, [12, 11, 12]) == [1, 4, 5]
    # G1Affine.Fold through the library's host helpers
    pts = [C.g1_mul(C.G1_GEN, rng.field(C.R)) for _ in range(3)] + [None]
    ch = rng.field(C.R)
    acc, coeff = np.zeros(8, dtype=np.uint64), 1
    for p in ocpu.g1_arr(pts):
        acc = lib.g1_add(acc, lib.g1_scalar_mul(p, bsb22.fr_to_limbs(coeff)))
        coeff = coeff * ch % C.R
    assert np.array_equal(acc, ocpu.g1_arr([CM.fold(pts, ch)])[0])
