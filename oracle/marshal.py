"""Proof.WriteTo / WriteRawTo from the oracle's integer proof (oracle; test infrastructure only): an independent
writer for the product's gnark_icicle_b200/marshal.py to be checked against (backend/groth16/bn254/marshal.go:30-70;
gnark-crypto point encodings restated from recollection -- self-consistency, not a pin).  BN254."""
import struct

from . import curve as C

P = C.P


def _largest(v):
    return v > (P - 1) // 2


def g1_bytes(pt, raw):
    if pt is None:
        return b"\x40" + bytes(63 if raw else 31)
    x, y = pt
    if raw:
        return x.to_bytes(32, "big") + y.to_bytes(32, "big")
    b = bytearray(x.to_bytes(32, "big"))
    b[0] |= 0xC0 if _largest(y) else 0x80
    return bytes(b)


def g2_bytes(pt, raw):
    if pt is None:
        return b"\x40" + bytes(127 if raw else 63)
    (x0, x1), (y0, y1) = pt
    if raw:
        return b"".join(v.to_bytes(32, "big") for v in (x1, x0, y1, y0))
    b = bytearray(x1.to_bytes(32, "big") + x0.to_bytes(32, "big"))
    b[0] |= 0xC0 if (_largest(y1) if y1 else _largest(y0)) else 0x80
    return bytes(b)


def proof_bytes(proof, raw=False):
    cms = proof.get("commitments", [])
    out = [g1_bytes(proof["ar"], raw), g2_bytes(proof["bs"], raw), g1_bytes(proof["krs"], raw), struct.pack(">I", len(cms))]
    out += [g1_bytes(c, raw) for c in cms]
    out.append(g1_bytes(proof.get("commitment_pok"), raw))
    return b"".join(out)


# ---- VerifyingKey / ProvingKey WriteTo / WriteRawTo (backend/groth16/bn254/marshal.go:99-166, :242-312) from the
# oracle's integer keys (oracle.groth16.setup); an independent writer for the product's to be compared with.
def _slice1(pts, raw):
    return struct.pack(">I", len(pts)) + b"".join(g1_bytes(p, raw) for p in pts)


def _slice2(pts, raw):
    return struct.pack(">I", len(pts)) + b"".join(g2_bytes(p, raw) for p in pts)


def _bool_bytes(bs):
    out = bytearray((len(bs) + 7) // 8)
    for i, b in enumerate(bs):
        if b:
            out[i >> 3] |= 1 << (i & 7)
    return struct.pack(">I", len(bs)) + bytes(out)


def vk_bytes(vk, raw=False):
    out = [g1_bytes(vk.g1_alpha, raw), g1_bytes(vk.g1_beta, raw), g2_bytes(vk.g2_beta, raw), g2_bytes(vk.g2_gamma, raw),
           g1_bytes(vk.g1_delta, raw), g2_bytes(vk.g2_delta, raw), _slice1(vk.g1_k, raw)]
    rows = vk.public_and_commitment_committed
    out.append(struct.pack(">I", len(rows)))
    for row in rows:
        out.append(struct.pack(">I", len(row)) + b"".join(struct.pack(">Q", v) for v in row))
    out.append(struct.pack(">I", len(vk.commitment_keys)))
    for k in vk.commitment_keys:
        out += [g2_bytes(k.g, raw), g2_bytes(k.g_sigma_neg, raw)]
    return b"".join(out)


def pk_bytes(pk, raw=False):
    from .curve import R, inv, FR_GEN
    d = pk.domain
    dom = struct.pack(">Q", d.n) + b"".join(v.to_bytes(32, "big") for v in (d.n_inv, d.gen, inv(d.gen, R), FR_GEN, inv(FR_GEN, R))) + b"\x00"
    out = [dom, g1_bytes(pk.g1_alpha, raw), g1_bytes(pk.g1_beta, raw), g1_bytes(pk.g1_delta, raw),
           _slice1(pk.g1_a, raw), _slice1(pk.g1_b, raw), _slice1(pk.g1_z, raw), _slice1(pk.g1_k, raw),
           g2_bytes(pk.g2_beta, raw), g2_bytes(pk.g2_delta, raw), _slice2(pk.g2_b, raw),
           struct.pack(">QQQ", len(pk.infinity_a), pk.nb_infinity_a, pk.nb_infinity_b),
           _bool_bytes(pk.infinity_a), _bool_bytes(pk.infinity_b), struct.pack(">I", len(pk.commitment_keys))]
    for k in pk.commitment_keys:
        out += [_slice1(k.basis, raw), _slice1(k.basis_exp_sigma, raw)]
    return b"".join(out)
