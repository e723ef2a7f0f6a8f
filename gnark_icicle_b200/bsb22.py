"""BSB22 commitments on the accelerated proving path.

Host side of backend/groth16/bn254/prove.go:81-139 (the icicle prover repeats the hint at icicle.go:152-190 but is
stale after it -- no proof of knowledge, SURVEY.md F5 -- so the CPU prover is the model):
the bsb22 solver-hint override that commits to the privately committed wires, hashes the commitment to the value
of the commitment wire, and -- after the solve -- the per-key proofs of knowledge folded with the "G16-BSB22"
challenge.  The two MSMs per commitment (pedersen Commit / ProveKnowledge over Basis / BasisExpSigma) run on the
device through ``g16_commit`` against resident commitment keys; hashing and folding are O(#commitments) host work,
exactly where the reference keeps them.

The Go shim calls gnark-crypto for hash_to_field and Marshal; this Python mirror carries its own small copies
(ExpandMsgXmd of RFC 9380 over SHA-256, fr.Hash, G1Affine.Marshal) because the product never touches the test-only CPU restatement.
"""
import hashlib
from dataclasses import dataclass, field as _field

import numpy as np

from . import lib as _lib

COMMITMENT_DST = b"bsb22-commitment"      # constraint/commitment.go:7
FOLD_DST = b"G16-BSB22"                   # prove.go:129

FR_MOD = {"bn254": 21888242871839275222246405745257275088548364400416034343698204186575808495617,
          "bls12_377": 8444461749428370424248824938781546531375899335154063827935233455917409239041}
FP_MOD = {"bn254": 21888242871839275222246405745257275088696311157297823662689037894645226208583,
          "bls12_377": 258664426012969094010652733694893533536393512754914660539884262666720468348340822774968888139573360124440321458177}


# ------------------------------------------------------------------ constraint.Groth16Commitment
@dataclass
class Groth16Commitment:
    """constraint/commitment.go:9-14."""
    public_and_commitment_committed: list = _field(default_factory=list)
    private_committed: list = _field(default_factory=list)
    commitment_index: int = 0
    nb_public_committed: int = 0


@dataclass
class CommitmentKey:
    """pedersen.ProvingKey (setup.go:57): two G1 affine arrays of equal length, gnark memory image."""
    basis: np.ndarray
    basis_exp_sigma: np.ndarray


def k_wire_indices(nb_public, nb_wires, info):
    """wire indices feeding the K MSM: W[nbPublic:] without privately committed and commitment wires
    (prove.go:241-246 filterHeap(wireValues[nbPublic:], nbPublic, ConcatAll(toRemove...)))."""
    remove = np.zeros(nb_wires, dtype=bool)
    for c in info:
        remove[np.asarray(c.private_committed, dtype=np.int64)] = True
        remove[c.commitment_index] = True
    idx = np.arange(nb_public, nb_wires, dtype=np.uint32)
    return idx[~remove[nb_public:]]


# ------------------------------------------------------------------ field / point (de)serialisation
def fr_to_limbs(v, curve="bn254"):
    """canonical int -> 4 x u64 Montgomery limbs (fr.Element memory image)."""
    m = (v % FR_MOD[curve]) * (1 << 256) % FR_MOD[curve]
    return np.array([(m >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)], dtype=np.uint64)


def fr_vec_to_limbs(vals, curve="bn254"):
    out = np.zeros((len(vals), 4), dtype=np.uint64)
    for i, v in enumerate(vals):
        out[i] = fr_to_limbs(int(v), curve)
    return out


def _fp_from_limbs(limbs, curve):
    p = FP_MOD[curve]
    v = 0
    for i, w in enumerate(limbs):
        v |= int(w) << (64 * i)
    return v * pow(1 << (64 * len(limbs)), -1, p) % p


def g1_marshal(pt, curve="bn254"):
    """G1Affine.Marshal: X || Y, big-endian canonical; the point at infinity is 0x40 followed by zeros."""
    pt = np.asarray(pt, dtype=np.uint64).reshape(-1)
    w = len(pt) // 2
    if not pt.any():
        return b"\x40" + bytes(16 * w - 1)
    return _fp_from_limbs(pt[:w], curve).to_bytes(8 * w, "big") + _fp_from_limbs(pt[w:], curve).to_bytes(8 * w, "big")


# ------------------------------------------------------------------ hash to field
def expand_msg_xmd(msg, dst, len_in_bytes):
    """RFC 9380 5.3.1, H = SHA-256 (gnark-crypto field/hash.ExpandMsgXmd)."""
    ell = (len_in_bytes + 31) // 32
    if ell > 255 or len(dst) > 255:
        raise ValueError("invalid lenInBytes / dst")
    dst_prime = dst + bytes([len(dst)])
    b0 = hashlib.sha256(bytes(64) + msg + len_in_bytes.to_bytes(2, "big") + b"\x00" + dst_prime).digest()
    blocks = [hashlib.sha256(b0 + b"\x01" + dst_prime).digest()]
    for i in range(2, ell + 1):
        mixed = bytes(x ^ y for x, y in zip(b0, blocks[-1]))
        blocks.append(hashlib.sha256(mixed + bytes([i]) + dst_prime).digest())
    return b"".join(blocks)[:len_in_bytes]


def fr_hash(msg, dst, count, curve="bn254"):
    """fr.Hash: 16 + 32 bytes per element, big-endian, reduced mod r."""
    L = 48
    buf = expand_msg_xmd(msg, dst, count * L)
    return [int.from_bytes(buf[i * L:(i + 1) * L], "big") % FR_MOD[curve] for i in range(count)]


class HashToField:
    """hash_to_field.New(dst) (backend.ProverConfig.HashToFieldFn default, prove.go:67-69)."""

    def __init__(self, dst=COMMITMENT_DST, curve="bn254"):
        self.dst, self.curve, self._buf = dst, curve, b""

    def write(self, data):
        self._buf += data

    def sum(self):
        return fr_hash(self._buf, self.dst, 1, self.curve)[0].to_bytes(32, "big")

    def reset(self):
        self._buf = b""

    def size(self):
        return 32


# ------------------------------------------------------------------ the prover's commitment session
class Session:
    """State of one Prove call: the hint override (prove.go:81-109) and the folded proof of knowledge
    (prove.go:117-135).  Commitment keys are uploaded once per proving key and stay resident."""

    def __init__(self, pk, info, ctx, hasher=None):
        if len(info) != len(pk.commitment_keys):
            raise _lib.G16Error(f"constraint system has {len(info)} commitments, the proving key {len(pk.commitment_keys)} keys")
        self.pk, self.info, self.ctx, self.curve = pk, info, ctx, pk.curve
        self.hasher = hasher or HashToField(curve=pk.curve)
        self.commitments = [None] * len(info)
        self.poks = [None] * len(info)
        if pk.commitment_device is None or pk.commitment_device[0] is not ctx:
            pk.commitment_device = (ctx, [ctx.commitment_key_upload(k.basis, k.basis_exp_sigma) for k in pk.commitment_keys])
        self.handles = pk.commitment_device[1]

    def hint(self, i, hashed, committed):
        """in[0] = i, in[1:] = public/commitment committed values then privately committed values (canonical ints);
        returns the value of commitment wire i."""
        c = self.info[i]
        if len(hashed) != len(c.public_and_commitment_committed) or len(committed) != len(c.private_committed):
            raise _lib.G16Error("bsb22 hint: wrong number of inputs")
        values = fr_vec_to_limbs(committed, self.curve)
        # Commit (prove.go:94) and, from the same upload, ProveKnowledge (prove.go:124)
        self.commitments[i], self.poks[i] = self.ctx.commit(self.handles[i], values, want_pok=True)
        fr_bytes = 32
        self.hasher.write(g1_marshal(self.commitments[i], self.curve) + b"".join(int(v).to_bytes(fr_bytes, "big") for v in hashed))
        out = self.hasher.sum()
        self.hasher.reset()
        return int.from_bytes(out[:min(fr_bytes, self.hasher.size())], "big") % FR_MOD[self.curve]

    def finish(self, commitment_wire_values):
        """(Commitments, CommitmentPok): challenge = fr.Hash(commitment wire values, "G16-BSB22", 1)[0],
        CommitmentPok = sum challenge^i poks[i] (G1Affine.Fold)."""
        if any(c is None for c in self.commitments):
            raise _lib.G16Error("the solver did not call the bsb22 hint for every commitment")
        w = len(self.pk.g1_alpha)
        if not self.info:
            return [], np.zeros(w, dtype=np.uint64)
        msg = b"".join(int(v).to_bytes(32, "big") for v in commitment_wire_values)
        challenge = fr_hash(msg, FOLD_DST, 1, self.curve)[0]
        acc, coeff = np.zeros(w, dtype=np.uint64), 1
        for pok in self.poks:
            term = pok if coeff == 1 else _lib.g1_scalar_mul(pok, fr_to_limbs(coeff, self.curve), curve=self.curve)
            acc = _lib.g1_add(acc, term, curve=self.curve)
            coeff = coeff * challenge % FR_MOD[self.curve]
        return list(self.commitments), acc
