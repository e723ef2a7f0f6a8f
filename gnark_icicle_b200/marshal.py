"""Proof.WriteTo / WriteRawTo / ReadFrom (backend/groth16/bn254/marshal.go:30-92) for the proofs this package
produces (wire formats).  Host-only code on Python ints; BN254.

Order on the wire: Ar, Bs, Krs, Commitments (u32 big-endian count, then the points), CommitmentPok -- compressed
by WriteTo, uncompressed by WriteRawTo; ReadFrom accepts either (the flag bits say which).  Point encodings are
gnark-crypto's (ecc/bn254/marshal.go, restated [recollection]; the reference holds no proof fixture): coordinates
big-endian canonical, E2 as A1 || A0, the two top bits of the first byte are the flag -- 00 uncompressed, 01
infinity (compressed or not), 10 / 11 compressed with the lexicographically smallest / largest y (y > (p-1)/2;
for E2 decided by A1 unless it is zero, then by A0).
"""
import struct

import numpy as np

from .bsb22 import FP_MOD

_P = FP_MOD["bn254"]
_RINV = pow(1 << 256, -1, _P)
_R = (1 << 256) % _P
_B1 = 3
_B2 = None   # 3 / (9 + u), filled lazily
M_MASK, M_UNCOMPRESSED, M_INFINITY, M_SMALLEST, M_LARGEST = 0xC0, 0x00, 0x40, 0x80, 0xC0


def _from_limbs(limbs):
    v = 0
    for i, w in enumerate(limbs):
        v |= int(w) << (64 * i)
    return v * _RINV % _P


def _to_limbs(v):
    m = v * _R % _P
    return [(m >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)]


def _coords(pt, n):
    pt = np.asarray(pt, dtype=np.uint64).reshape(-1)
    assert pt.size == 4 * n
    return None if not pt.any() else [_from_limbs(pt[4 * i:4 * i + 4]) for i in range(n)]


def _lex_largest(v):
    return v > (_P - 1) // 2


def _be(v):
    return v.to_bytes(32, "big")


def g1_encode(pt, raw=False):
    c = _coords(pt, 2)
    if c is None:
        return bytes([M_INFINITY]) + bytes(63 if raw else 31)
    if raw:
        return _be(c[0]) + _be(c[1])
    out = bytearray(_be(c[0]))
    out[0] |= M_LARGEST if _lex_largest(c[1]) else M_SMALLEST
    return bytes(out)


def g2_encode(pt, raw=False):
    c = _coords(pt, 4)          # X.A0 X.A1 Y.A0 Y.A1
    if c is None:
        return bytes([M_INFINITY]) + bytes(127 if raw else 63)
    if raw:
        return _be(c[1]) + _be(c[0]) + _be(c[3]) + _be(c[2])
    out = bytearray(_be(c[1]) + _be(c[0]))
    largest = _lex_largest(c[3]) if c[3] else _lex_largest(c[2])
    out[0] |= M_LARGEST if largest else M_SMALLEST
    return bytes(out)


# ---- decompression: square roots in Fp (p = 3 mod 4) and Fp2 = Fp[u]/(u^2+1)
def _fp_sqrt(a):
    r = pow(a, (_P + 1) // 4, _P)
    return r if r * r % _P == a % _P else None


def _f2_mul(a, b):
    return ((a[0] * b[0] - a[1] * b[1]) % _P, (a[0] * b[1] + a[1] * b[0]) % _P)


def _f2_sqrt(a):
    a0, a1 = a
    if a1 == 0:
        r = _fp_sqrt(a0)
        if r is not None:
            return (r, 0)
        r = _fp_sqrt(-a0 % _P)
        return None if r is None else (0, r)
    s = _fp_sqrt((a0 * a0 + a1 * a1) % _P)          # norm
    if s is None:
        return None
    half = pow(2, -1, _P)
    x0 = _fp_sqrt((a0 + s) * half % _P)
    if x0 is None:
        x0 = _fp_sqrt((a0 - s) * half % _P)
    if x0 is None or x0 == 0:
        return None
    x1 = a1 * pow(2 * x0, -1, _P) % _P
    return (x0, x1) if _f2_mul((x0, x1), (x0, x1)) == (a0 % _P, a1 % _P) else None


def _twist_b():
    global _B2
    if _B2 is None:
        d = pow(9 * 9 + 1, -1, _P)                  # 1 / (9 + u) = (9 - u) / 82
        _B2 = (3 * 9 * d % _P, -3 * d % _P)
    return _B2


def g1_decode(buf, off=0):
    """(limb array, bytes consumed)."""
    flag = buf[off] & M_MASK
    if flag == M_INFINITY:
        raise ValueError("g1_decode: use proof_read_from (the length of an encoded infinity depends on the encoding in use)")
    if flag == M_UNCOMPRESSED:
        x, y = int.from_bytes(buf[off:off + 32], "big"), int.from_bytes(buf[off + 32:off + 64], "big")
        n = 64
    else:
        x = int.from_bytes(bytes([buf[off] & ~M_MASK & 0xFF]) + buf[off + 1:off + 32], "big")
        y = _fp_sqrt((x * x * x + _B1) % _P)
        if y is None:
            raise ValueError("invalid compressed G1 point")
        if _lex_largest(y) != (flag == M_LARGEST):
            y = _P - y
        n = 32
    if x >= _P or y >= _P or (y * y - x * x * x - _B1) % _P:
        raise ValueError("G1 point is not on the curve")
    return np.array(_to_limbs(x) + _to_limbs(y), dtype=np.uint64), n


def g2_decode(buf, off=0):
    flag = buf[off] & M_MASK
    if flag == M_INFINITY:
        raise ValueError("g2_decode: use proof_read_from")
    if flag == M_UNCOMPRESSED:
        x1, x0, y1, y0 = (int.from_bytes(buf[off + 32 * i:off + 32 * i + 32], "big") for i in range(4))
        n = 128
    else:
        x1 = int.from_bytes(bytes([buf[off] & ~M_MASK & 0xFF]) + buf[off + 1:off + 32], "big")
        x0 = int.from_bytes(buf[off + 32:off + 64], "big")
        x = (x0, x1)
        b = _twist_b()
        x3 = _f2_mul(_f2_mul(x, x), x)
        y = _f2_sqrt(((x3[0] + b[0]) % _P, (x3[1] + b[1]) % _P))
        if y is None:
            raise ValueError("invalid compressed G2 point")
        largest = _lex_largest(y[1]) if y[1] else _lex_largest(y[0])
        if largest != (flag == M_LARGEST):
            y = (-y[0] % _P, -y[1] % _P)
        y0, y1 = y
        n = 64
    b = _twist_b()
    x = (x0, x1)
    x3 = _f2_mul(_f2_mul(x, x), x)
    if _f2_mul((y0, y1), (y0, y1)) != ((x3[0] + b[0]) % _P, (x3[1] + b[1]) % _P):
        raise ValueError("G2 point is not on the curve")
    return np.array(_to_limbs(x0) + _to_limbs(x1) + _to_limbs(y0) + _to_limbs(y1), dtype=np.uint64), n


def proof_write_to(proof, raw=False):
    """Proof.WriteTo (compressed) / WriteRawTo (raw=True)."""
    if len(np.asarray(proof.Ar).reshape(-1)) != 8:
        raise ValueError("proof wire format is implemented for BN254")
    out = [g1_encode(proof.Ar, raw), g2_encode(proof.Bs, raw), g1_encode(proof.Krs, raw),
           struct.pack(">I", len(proof.Commitments))]
    out += [g1_encode(c, raw) for c in proof.Commitments]
    pok = proof.CommitmentPok if proof.CommitmentPok is not None else np.zeros(8, dtype=np.uint64)
    out.append(g1_encode(pok, raw))
    return b"".join(out)


def proof_read_from(buf):
    """Proof.ReadFrom: either encoding; returns (Proof, bytes read)."""
    from .groth16 import Proof
    # the encoding is uniform over a proof: Ar is never the point at infinity in a valid proof, so its flag tells
    raw = (buf[0] & M_MASK) == M_UNCOMPRESSED
    off = 0

    def g1():
        nonlocal off
        if (buf[off] & M_MASK) == M_INFINITY:
            off += 64 if raw else 32
            return np.zeros(8, dtype=np.uint64)
        pt, n = g1_decode(buf, off)
        off += n
        return pt

    def g2():
        nonlocal off
        if (buf[off] & M_MASK) == M_INFINITY:
            off += 128 if raw else 64
            return np.zeros(16, dtype=np.uint64)
        pt, n = g2_decode(buf, off)
        off += n
        return pt

    ar, bs, krs = g1(), g2(), g1()
    (count,) = struct.unpack(">I", buf[off:off + 4])
    off += 4
    commitments = [g1() for _ in range(count)]
    pok = g1()
    return Proof(Ar=ar, Bs=bs, Krs=krs, Commitments=commitments, CommitmentPok=pok), off


# ---------------------------------------------------------------------------------------------------------------------
# VerifyingKey / ProvingKey WriteTo, WriteRawTo, ReadFrom (backend/groth16/bn254/marshal.go:99-236, :242-385).
# Container conventions of gnark-crypto's curve.Encoder, restated [recollection] like the point encodings above:
# uint32 / uint64 big-endian; a point slice is a uint32 count followed by the points; [][]uint64 is a uint32 count of
# rows, each row a uint32 count followed by its uint64s; []bool is a uint32 count followed by the bits packed LSB-first
# (the convention dump.py uses for the same fields); fft.Domain.WriteTo as in dump.py; pedersen.VerifyingKey = G,
# GSigmaNeg (G2), pedersen.ProvingKey = Basis, BasisExpSigma (G1 slices).
from dataclasses import dataclass, field as _field


@dataclass
class PedersenVerifyingKey:
    """pedersen.VerifyingKey (verify.go:107-115 consumes it)."""
    G: np.ndarray            # (16,) G2Affine
    GSigmaNeg: np.ndarray    # (16,)


@dataclass
class VerifyingKey:
    """setup.go:60-83 fields that go on the wire (vk.e and the negated points are recomputed by Precompute)."""
    g1_alpha: np.ndarray
    g1_beta: np.ndarray
    g2_beta: np.ndarray
    g2_gamma: np.ndarray
    g1_delta: np.ndarray
    g2_delta: np.ndarray
    g1_k: np.ndarray                                     # (nbPublic + nbCommitments, 8)
    public_and_commitment_committed: list = _field(default_factory=list)
    commitment_keys: list = _field(default_factory=list)   # []PedersenVerifyingKey


def _g1_slice(arr, raw):
    arr = np.asarray(arr, dtype=np.uint64).reshape(-1, 8)
    return struct.pack(">I", len(arr)) + b"".join(g1_encode(p, raw) for p in arr)


def _g2_slice(arr, raw):
    arr = np.asarray(arr, dtype=np.uint64).reshape(-1, 16)
    return struct.pack(">I", len(arr)) + b"".join(g2_encode(p, raw) for p in arr)


def _bools(b):
    b = np.asarray(b, dtype=np.uint8)
    return struct.pack(">I", len(b)) + np.packbits(b, bitorder="little").tobytes()


def vk_write_to(vk, raw=False):
    """VerifyingKey.WriteTo (compressed) / WriteRawTo: [alpha]1,[beta]1,[beta]2,[gamma]2,[delta]1,[delta]2,
    uint32(len(Kvk)),[Kvk]1, PublicAndCommitmentCommitted, uint32(len(CommitmentKeys)), the Pedersen keys."""
    out = [g1_encode(vk.g1_alpha, raw), g1_encode(vk.g1_beta, raw), g2_encode(vk.g2_beta, raw), g2_encode(vk.g2_gamma, raw),
           g1_encode(vk.g1_delta, raw), g2_encode(vk.g2_delta, raw), _g1_slice(vk.g1_k, raw)]
    rows = vk.public_and_commitment_committed or []
    out.append(struct.pack(">I", len(rows)))
    for row in rows:
        out.append(struct.pack(">I", len(row)) + b"".join(struct.pack(">Q", int(v)) for v in row))
    out.append(struct.pack(">I", len(vk.commitment_keys)))
    for k in vk.commitment_keys:
        out += [g2_encode(k.G, raw), g2_encode(k.GSigmaNeg, raw)]
    return b"".join(out)


class _Reader:
    def __init__(self, buf, off=0):
        self.buf, self.off = buf, off
        self.raw = None            # decided by the first point met

    def u32(self):
        (v,) = struct.unpack(">I", self.buf[self.off:self.off + 4])
        self.off += 4
        return v

    def u64(self):
        (v,) = struct.unpack(">Q", self.buf[self.off:self.off + 8])
        self.off += 8
        return v

    def _point(self, words, decode):
        flag = self.buf[self.off] & M_MASK
        if self.raw is None:
            # uniform over a key; an infinity flag says nothing, so a leading infinity leaves the question open
            if flag != M_INFINITY:
                self.raw = flag == M_UNCOMPRESSED
        if flag == M_INFINITY:
            if self.raw is None:
                raise ValueError("cannot tell the encoding from a leading point at infinity")
            self.off += (8 * words) if self.raw else (4 * words)
            return np.zeros(words, dtype=np.uint64)
        pt, n = decode(self.buf, self.off)
        self.off += n
        return pt

    def g1(self):
        return self._point(8, g1_decode)

    def g2(self):
        return self._point(16, g2_decode)

    def g1_slice(self):
        n = self.u32()
        return np.array([self.g1() for _ in range(n)], dtype=np.uint64).reshape(n, 8)

    def g2_slice(self):
        n = self.u32()
        return np.array([self.g2() for _ in range(n)], dtype=np.uint64).reshape(n, 16)

    def bools(self):
        n = self.u32()
        rawb = np.frombuffer(self.buf[self.off:self.off + (n + 7) // 8], dtype=np.uint8)
        self.off += (n + 7) // 8
        return np.unpackbits(rawb, bitorder="little")[:n].astype(np.uint8)


def vk_read_from(buf):
    """VerifyingKey.ReadFrom: either encoding; returns (VerifyingKey, bytes read)."""
    r = _Reader(buf)
    alpha, beta1, beta2, gamma2, delta1, delta2 = r.g1(), r.g1(), r.g2(), r.g2(), r.g1(), r.g2()
    k = r.g1_slice()
    rows = [[r.u64() for _ in range(r.u32())] for _ in range(r.u32())]
    keys = [PedersenVerifyingKey(r.g2(), r.g2()) for _ in range(r.u32())]
    return VerifyingKey(alpha, beta1, beta2, gamma2, delta1, delta2, k, rows, keys), r.off


def pk_write_to(pk, raw=False):
    """ProvingKey.WriteTo / WriteRawTo (marshal.go:242-312): Domain, [alpha]1,[beta]1,[delta]1, A, B, Z, K, [beta]2,
    [delta]2, B2, nbWires, NbInfinityA, NbInfinityB, InfinityA, InfinityB, uint32(len(CommitmentKeys)), the keys."""
    from .dump import _domain_bytes
    if pk.curve != "bn254":
        raise ValueError("key wire format is implemented for BN254")
    ia, ib = np.asarray(pk.infinity_a, dtype=np.uint8), np.asarray(pk.infinity_b, dtype=np.uint8)
    out = [_domain_bytes(pk.domain_n, pk.curve),
           g1_encode(pk.g1_alpha, raw), g1_encode(pk.g1_beta, raw), g1_encode(pk.g1_delta, raw),
           _g1_slice(pk.g1_a, raw), _g1_slice(pk.g1_b, raw), _g1_slice(pk.g1_z, raw), _g1_slice(pk.g1_k, raw),
           g2_encode(pk.g2_beta, raw), g2_encode(pk.g2_delta, raw), _g2_slice(pk.g2_b, raw),
           struct.pack(">QQQ", len(ia), int(ia.sum()), int(ib.sum())), _bools(ia), _bools(ib),
           struct.pack(">I", len(pk.commitment_keys))]
    for k in pk.commitment_keys:
        out += [_g1_slice(k.basis, raw), _g1_slice(k.basis_exp_sigma, raw)]
    return b"".join(out)


def pk_read_from(buf, nb_public=0):
    """ProvingKey.ReadFrom: either encoding; returns (gnark_icicle_b200.ProvingKey, bytes read).  nb_public is not on
    the wire (the reference takes it from the constraint system at prove time)."""
    from .bsb22 import CommitmentKey
    from .dump import _domain_bytes
    from .groth16 import ProvingKey
    (n,) = struct.unpack(">Q", buf[:8])
    dom = _domain_bytes(n, "bn254")
    if bytes(buf[:len(dom)]) != dom:
        raise ValueError("unexpected fft.Domain encoding")
    r = _Reader(buf, len(dom))
    alpha, beta, delta = r.g1(), r.g1(), r.g1()
    a, b, z, k = r.g1_slice(), r.g1_slice(), r.g1_slice(), r.g1_slice()
    beta2, delta2 = r.g2(), r.g2()
    b2 = r.g2_slice()
    nb_wires, nia, nib = r.u64(), r.u64(), r.u64()
    ia, ib = r.bools(), r.bools()
    if len(ia) != nb_wires or len(ib) != nb_wires or int(ia.sum()) != nia or int(ib.sum()) != nib:
        raise ValueError("infinity masks do not match their counts")
    keys = [CommitmentKey(r.g1_slice(), r.g1_slice()) for _ in range(r.u32())]
    pk = ProvingKey(domain_n=n, g1_alpha=alpha, g1_beta=beta, g1_delta=delta, g1_a=a, g1_b=b, g1_z=z, g1_k=k,
                    g2_beta=beta2, g2_delta=delta2, g2_b=b2, infinity_a=ia, infinity_b=ib, nb_wires=int(nb_wires),
                    nb_public=nb_public, commitment_keys=keys)
    return pk, r.off
