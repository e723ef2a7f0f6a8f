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
