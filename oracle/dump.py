"""ProvingKey.WriteDump image from the oracle's big-int proving key (oracle; test infrastructure only).

Independent of gnark_icicle_b200/dump.py (it works from Python ints, not from limb arrays) so that the product's
dump reader / device loader has something other than its own writer to be checked against.  Layout:
backend/groth16/bn254/marshal.go:386-456; gnark-crypto encodings restated from recollection (see the product module's
header) -- the reference holds no dump fixture, so this is self-consistency, not a pin.
"""
import struct

from . import curve as C
from .curve import R, inv


def _g1_raw(pt):
    if pt is None:
        return b"\x40" + bytes(C.G1_BYTES - 1)
    return pt[0].to_bytes(C.FP_BYTES, "big") + pt[1].to_bytes(C.FP_BYTES, "big")


def _g2_raw(pt):
    if pt is None:
        return b"\x40" + bytes(C.G2_BYTES - 1)
    (x0, x1), (y0, y1) = pt
    return b"".join(v.to_bytes(C.FP_BYTES, "big") for v in (x1, x0, y1, y0))


def _bools(bs):
    out = bytearray((len(bs) + 7) // 8)
    for i, v in enumerate(bs):
        if v:
            out[i >> 3] |= 1 << (i & 7)
    return struct.pack(">I", len(bs)) + bytes(out)


def _slice(raw, count):
    return struct.pack("<Q", count) + raw


def write_dump(pk):
    d = pk.domain
    out = [struct.pack("<Q", 0xDEADBEEF), struct.pack(">Q", d.n)]
    g = C.FR_GEN
    out += [v.to_bytes(32, "big") for v in (d.n_inv, d.gen, inv(d.gen, R), g, inv(g, R))] + [b"\x00"]
    out += [_g1_raw(pk.g1_alpha), _g1_raw(pk.g1_beta), _g1_raw(pk.g1_delta), _g2_raw(pk.g2_beta), _g2_raw(pk.g2_delta)]
    out.append(struct.pack(">QQQ", len(pk.infinity_a), pk.nb_infinity_a, pk.nb_infinity_b))
    out += [_bools(pk.infinity_a), _bools(pk.infinity_b)]
    keys = getattr(pk, "commitment_keys", [])
    out.append(struct.pack(">I", len(keys)))
    for pts in (pk.g1_a, pk.g1_b, pk.g1_z, pk.g1_k):
        out.append(_slice(C.g1_vec_to_bytes(pts), len(pts)))
    out.append(_slice(C.g2_vec_to_bytes(pk.g2_b), len(pk.g2_b)))
    for k in keys:
        out.append(_slice(C.g1_vec_to_bytes(k.basis), len(k.basis)))
        out.append(_slice(C.g1_vec_to_bytes(k.basis_exp_sigma), len(k.basis_exp_sigma)))
    return b"".join(out)
