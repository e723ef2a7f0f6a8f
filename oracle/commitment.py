"""BSB22 commitments of the Groth16 prover on Python ints (oracle; test infrastructure only).

Restates what backend/groth16/bn254/prove.go:81-139, setup.go:278-307 and verify.go:82-126 take from gnark-crypto
(v0.14.1-0.20241010154951-6638408a49f3, not vendored -- written from its published algorithms):

  * fr.Hash / ecc/bn254/fr/hash_to_field: ExpandMsgXmd of RFC 9380 section 5.3.1 over SHA-256, 16 + 32 bytes per
    element, big-endian, reduced mod r; the hash.Hash wrapper `hash_to_field.New(dst)` whose Sum is
    fr.Hash(written bytes, dst, 1)[0].Bytes().  Pinned by the RFC's own expand_message_xmd vectors
    (tests/test_oracle_commitment.py).
  * G1Affine.Marshal (uncompressed: X || Y big-endian canonical, infinity = 0x40 then zeros).
  * ecc/bn254/fr/pedersen: Setup (BasisExpSigma = sigma * Basis, vk = {G, GSigmaNeg = -sigma * G}), Commit and
    ProveKnowledge (MSMs over Basis / BasisExpSigma), G1Affine.Fold (sum r^i P_i), BatchVerifyMultiVk.  The pairing
    equation e(commitment, GSigmaNeg) e(pok, G) == 1 is the one the reference's Solidity verifier hard-codes
    (backend/groth16/bn254/solidity.go:580-598).
  * constraint.SerializeCommitment / Groth16Commitments (constraint/commitment.go:9-14,43-85).
"""
import hashlib
from dataclasses import dataclass, field as _field

from . import curve as C
from .curve import R

COMMITMENT_DST = b"bsb22-commitment"      # constraint/commitment.go:7
FOLD_DST = b"G16-BSB22"                   # prove.go:129, verify.go:108


# ------------------------------------------------------------------ hash to field
def expand_msg_xmd(msg, dst, len_in_bytes):
    """RFC 9380 5.3.1 with H = SHA-256 (b_in_bytes 32, s_in_bytes 64)."""
    ell = (len_in_bytes + 31) // 32
    if ell > 255 or len(dst) > 255:
        raise ValueError("invalid lenInBytes / dst")
    dst_prime = dst + bytes([len(dst)])
    b0 = hashlib.sha256(bytes(64) + msg + len_in_bytes.to_bytes(2, "big") + b"\x00" + dst_prime).digest()
    b = hashlib.sha256(b0 + b"\x01" + dst_prime).digest()
    out = b
    for i in range(2, ell + 1):
        b = hashlib.sha256(bytes(x ^ y for x, y in zip(b0, b)) + bytes([i]) + dst_prime).digest()
        out += b
    return out[:len_in_bytes]


def fr_hash(msg, dst, count):
    """fr.Hash: L = 16 + fr.Bytes pseudo-random bytes per element, big-endian, mod r."""
    L = 16 + C.FR_BYTES
    buf = expand_msg_xmd(msg, dst, count * L)
    return [int.from_bytes(buf[i * L:(i + 1) * L], "big") % R for i in range(count)]


class HashToField:
    """hash_to_field.New(dst): Write appends, Sum = fr.Hash(buffer, dst, 1)[0] as 32 big-endian bytes."""

    def __init__(self, dst=COMMITMENT_DST):
        self.dst, self.buf = dst, b""

    def write(self, data):
        self.buf += data

    def sum(self):
        return fr_hash(self.buf, self.dst, 1)[0].to_bytes(C.FR_BYTES, "big")

    def reset(self):
        self.buf = b""

    def size(self):
        return C.FR_BYTES


def fr_marshal(v):
    return (v % R).to_bytes(C.FR_BYTES, "big")


def g1_marshal(pt):
    """G1Affine.Marshal = RawBytes: uncompressed, big-endian canonical coordinates; infinity flag 0b01 in the top bits."""
    if pt is None:
        return b"\x40" + bytes(C.G1_BYTES - 1)
    return pt[0].to_bytes(C.FP_BYTES, "big") + pt[1].to_bytes(C.FP_BYTES, "big")


def serialize_commitment(commitment_bytes, public_committed, field_byte_len):
    """constraint/commitment.go:76-89."""
    return commitment_bytes + b"".join(int(v).to_bytes(field_byte_len, "big") for v in public_committed)


def solve_commitment_wire(commitment_pt, hashed_values, hasher=None):
    """the bsb22 hint body (prove.go:94-108) / verify.go:89-104: H(commitment || public committed values) -> Fr."""
    h = hasher or HashToField()
    h.write(serialize_commitment(g1_marshal(commitment_pt), hashed_values, (R.bit_length() - 1) // 8 + 1))
    out = h.sum()
    h.reset()
    nb = min(C.FR_BYTES, h.size())
    return int.from_bytes(out[:nb], "big") % R       # fr.Element.SetBytes reduces


# ------------------------------------------------------------------ constraint.Groth16Commitment
@dataclass
class Groth16Commitment:
    """constraint/commitment.go:9-14."""
    public_and_commitment_committed: list = _field(default_factory=list)
    private_committed: list = _field(default_factory=list)
    commitment_index: int = 0
    nb_public_committed: int = 0


def commitment_indexes(info):
    return [c.commitment_index for c in info]


def get_public_and_commitment_committed(info, translation, offset):
    """constraint/commitment.go:53-74: commitment-committed wires are renumbered to their position among the
    commitment wires plus `offset` (= number of public wires), i.e. to where the verifier appends them."""
    res = []
    for c in info:
        row = list(c.public_and_commitment_committed[:c.nb_public_committed])
        k = 0
        for wire in c.public_and_commitment_committed[c.nb_public_committed:]:
            while translation[k] != wire:
                k += 1
            row.append(k + offset)
        res.append(row)
    return res


# ------------------------------------------------------------------ pedersen
@dataclass
class PedersenPK:
    basis: list
    basis_exp_sigma: list


@dataclass
class PedersenVK:
    g: tuple
    g_sigma_neg: tuple


def pedersen_setup(basis, g2_point, sigma):
    """pedersen.Setup(bases[i:i+1], WithG2Point(cG2)) with the secret sigma injected (setup.go:298-305)."""
    pk = PedersenPK(list(basis), [C.g1_mul(b, sigma) for b in basis])
    vk = PedersenVK(g2_point, C.g2_mul(g2_point, (-sigma) % R))
    return pk, vk


def pedersen_commit(points, values, msm=None):
    """ProvingKey.Commit (Basis) / ProveKnowledge (BasisExpSigma): a MultiExp."""
    if len(points) != len(values):
        raise ValueError("must have as many values as basis elements")
    if msm:
        return msm(points, values)
    acc = None
    for p, v in zip(points, values):
        acc = C.g1_add(acc, C.g1_mul(p, v % R))
    return acc


def fold(points, challenge):
    """G1Affine.Fold: sum challenge^i * points[i] (prove.go:133)."""
    acc, coeff = None, 1
    for p in points:
        acc = C.g1_add(acc, C.g1_mul(p, coeff))
        coeff = coeff * challenge % R
    return acc


def fold_challenge(commitment_wire_values):
    """prove.go:125-129 / verify.go:105-108."""
    return fr_hash(b"".join(fr_marshal(v) for v in commitment_wire_values), FOLD_DST, 1)[0]


def batch_verify_multi_vk(vks, commitments, pok, challenge, pairing_product):
    """pedersen.BatchVerifyMultiVk with one folded proof: prod e(r^i C_i, GSigmaNeg_i) * e(pok, G) == 1."""
    if len(vks) != len(commitments):
        raise ValueError("commitments / verifying keys length mismatch")
    for v in vks:
        if v.g != vks[0].g:
            raise ValueError("parameter mismatch: G2 element")
    pairs, coeff = [], 1
    for v, c in zip(vks, commitments):
        if c is not None:
            pairs.append((C.g1_mul(c, coeff), v.g_sigma_neg))
        coeff = coeff * challenge % R
    if pok is not None:
        pairs.append((pok, vks[0].g))
    from .pairing import F12_ONE
    return pairing_product(pairs) == F12_ONE if pairs else True
