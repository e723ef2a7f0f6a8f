"""ProvingKey.WriteDump / ReadDump (backend/groth16/bn254/marshal.go:386-550) and the direct device load built on
it.

A dump is: the gnark-crypto unsafe marker, the fft.Domain, a small header in the curve's raw encoding (alpha, beta,
delta, wire counts, InfinityA/B, number of commitment keys), then the point slices G1.A, G1.B, G1.Z, G1.K, G2.B and
every commitment key's Basis / BasisExpSigma written by unsafe.WriteSlice -- a u64 length followed by the slice's raw
memory.  That raw memory (little-endian Montgomery limbs, affine, (0,0) = infinity) is exactly what the device keeps,
so `read_dump_to_device` parses the header, notes where each slice starts, and lets libg16b200 stream the slices
file -> pinned staging -> HBM (`g16_pk_upload_from_file`) without ever materialising them as host arrays.

The header encodings below are those of gnark-crypto v0.14.1-0.20241010154951-6638408a49f3 (not vendored in the
reference; restated [recollection]): utils/unsafe marker 0xdeadbeef as a native u64; fft.Domain.WriteTo = Cardinality
(u64 big-endian), CardinalityInv, Generator, GeneratorInv, FrMultiplicativeGen, FrMultiplicativeGenInv (32-byte
big-endian canonical), withPrecompute (1 byte); Encoder(RawEncoding): G1 = X || Y, G2 = X.A1 || X.A0 || Y.A1 || Y.A0
big-endian canonical uncompressed (infinity: 0x40 then zeros), integers big-endian, []bool = u32 length then bits
packed LSB-first.  In the Go shim the header is decoded by gnark itself, so only this Python mirror depends on them.
"""
import io
import os
import struct
from dataclasses import dataclass, field as _field

import numpy as np

from . import lib as _lib
from .bsb22 import FR_MOD, FP_MOD, CommitmentKey
from .groth16 import ProvingKey, ProverConfig

MARKER = 0xDEADBEEF
_COSET_GEN = {"bn254": 5, "bls12_377": 22}
_TWO_ADICITY = {"bn254": 28, "bls12_377": 47}


def _fp_words(curve):
    return _lib.CURVES[curve][1] // 2


def _fp_from_mont(limbs, curve):
    p = FP_MOD[curve]
    v = 0
    for i, w in enumerate(limbs):
        v |= int(w) << (64 * i)
    return v * pow(1 << (64 * len(limbs)), -1, p) % p


def _fp_to_mont(v, curve):
    w = _fp_words(curve)
    m = v * (1 << (64 * w)) % FP_MOD[curve]
    return [(m >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(w)]


def _point_raw(pt, curve, g2):
    """uncompressed raw encoding of one affine point given as Montgomery limbs."""
    w = _fp_words(curve)
    pt = np.asarray(pt, dtype=np.uint64).reshape(-1)
    n = 4 if g2 else 2
    if not pt.any():
        return b"\x40" + bytes(n * 8 * w - 1)
    c = [_fp_from_mont(pt[i * w:(i + 1) * w], curve).to_bytes(8 * w, "big") for i in range(n)]
    return c[0] + c[1] if not g2 else c[1] + c[0] + c[3] + c[2]      # E2 is written A1 first


def _point_from_raw(buf, curve, g2):
    w = _fp_words(curve)
    n = 4 if g2 else 2
    if buf[0] & 0xC0 == 0x40:
        return np.zeros(n * w, dtype=np.uint64)
    c = [int.from_bytes(buf[i * 8 * w:(i + 1) * 8 * w], "big") for i in range(n)]
    if g2:
        c = [c[1], c[0], c[3], c[2]]
    return np.array([x for v in c for x in _fp_to_mont(v, curve)], dtype=np.uint64)


def _pack_bools(b):
    b = np.asarray(b, dtype=np.uint8)
    return struct.pack(">I", len(b)) + np.packbits(b, bitorder="little").tobytes()


def _read_bools(f):
    (n,) = struct.unpack(">I", f.read(4))
    raw = np.frombuffer(f.read((n + 7) // 8), dtype=np.uint8)
    return np.unpackbits(raw, bitorder="little")[:n].astype(np.uint8)


def _domain_bytes(n, curve):
    r = FR_MOD[curve]
    root = pow(_COSET_GEN[curve], (r - 1) >> _TWO_ADICITY[curve], r)
    gen = pow(root, (1 << _TWO_ADICITY[curve]) // n, r)
    g = _COSET_GEN[curve]
    vals = [pow(n, -1, r), gen, pow(gen, -1, r), g, pow(g, -1, r)]
    return struct.pack(">Q", n) + b"".join(v.to_bytes(32, "big") for v in vals) + b"\x00"


def _write_slice(f, arr):
    arr = np.ascontiguousarray(arr, dtype=np.uint64)
    f.write(struct.pack("<Q", arr.shape[0] if arr.ndim == 2 else 0))
    f.write(arr.tobytes())


def write_dump(pk, f):
    """ProvingKey.WriteDump (marshal.go:386-456)."""
    c = pk.curve
    f.write(struct.pack("<Q", MARKER))
    f.write(_domain_bytes(pk.domain_n, c))
    for pt, g2 in ((pk.g1_alpha, 0), (pk.g1_beta, 0), (pk.g1_delta, 0), (pk.g2_beta, 1), (pk.g2_delta, 1)):
        f.write(_point_raw(pt, c, g2))
    ia, ib = np.asarray(pk.infinity_a, dtype=np.uint8), np.asarray(pk.infinity_b, dtype=np.uint8)
    f.write(struct.pack(">QQQ", len(ia), int(ia.sum()), int(ib.sum())))
    f.write(_pack_bools(ia))
    f.write(_pack_bools(ib))
    f.write(struct.pack(">I", len(pk.commitment_keys)))
    for arr in (pk.g1_a, pk.g1_b, pk.g1_z, pk.g1_k, pk.g2_b):
        _write_slice(f, arr)
    for k in pk.commitment_keys:
        _write_slice(f, k.basis)
        _write_slice(f, k.basis_exp_sigma)


@dataclass
class DumpLayout:
    """what the header of a dump says, and where each raw slice lives: name -> (byte offset of element 0, count)."""
    domain_n: int
    points: dict
    nb_wires: int
    infinity_a: np.ndarray
    infinity_b: np.ndarray
    spans: dict
    commitment_spans: list = _field(default_factory=list)     # [(basis span, basis_exp_sigma span)]


def read_dump_layout(f, curve="bn254"):
    """ReadDump up to the first slice (marshal.go:458-509), then only the slice headers."""
    g1w, g2w = _lib.CURVES[curve][1], _lib.CURVES[curve][2]
    (marker,) = struct.unpack("<Q", f.read(8))
    if marker != MARKER:
        raise _lib.G16Error("read marker: not a gnark-crypto unsafe dump")
    (n,) = struct.unpack(">Q", f.read(8))
    f.read(5 * 32 + 1)                                   # the domain's derived fields are recomputed on the device
    pts = {}
    for name, g2 in (("g1_alpha", 0), ("g1_beta", 0), ("g1_delta", 0), ("g2_beta", 1), ("g2_delta", 1)):
        pts[name] = _point_from_raw(f.read((g2w if g2 else g1w) * 8), curve, g2)
    nb_wires, nb_inf_a, nb_inf_b = struct.unpack(">QQQ", f.read(24))
    ia, ib = _read_bools(f), _read_bools(f)
    if len(ia) != nb_wires or len(ib) != nb_wires or int(ia.sum()) != nb_inf_a or int(ib.sum()) != nb_inf_b:
        raise _lib.G16Error("dump header is inconsistent (InfinityA/B against nbWires / NbInfinityA/B)")
    (nb_commitments,) = struct.unpack(">I", f.read(4))

    def span(words):
        (count,) = struct.unpack("<Q", f.read(8))
        off = f.tell()
        f.seek(count * words * 8, os.SEEK_CUR)
        return off, count

    spans = {}
    for name, words in (("g1_a", g1w), ("g1_b", g1w), ("g1_z", g1w), ("g1_k", g1w), ("g2_b", g2w)):
        spans[name] = span(words)
    cspans = [(span(g1w), span(g1w)) for _ in range(nb_commitments)]
    end = f.tell()
    f.seek(0, os.SEEK_END)
    if f.tell() < end:
        raise _lib.G16Error("dump is truncated")
    return DumpLayout(n, pts, nb_wires, ia, ib, spans, cspans)


def _read_span(f, span, words):
    off, count = span
    f.seek(off)
    return np.frombuffer(f.read(count * words * 8), dtype=np.uint64).reshape(count, words).copy()


def read_dump(path, curve="bn254", nb_public=0):
    """ProvingKey.ReadDump (marshal.go:458-550): everything into host arrays, like the reference."""
    g1w, g2w = _lib.CURVES[curve][1], _lib.CURVES[curve][2]
    with open(path, "rb") as f:
        lay = read_dump_layout(f, curve)
        arrs = {k: _read_span(f, lay.spans[k], g2w if k == "g2_b" else g1w) for k in lay.spans}
        cks = [CommitmentKey(_read_span(f, b, g1w), _read_span(f, e, g1w)) for b, e in lay.commitment_spans]
    return ProvingKey(domain_n=lay.domain_n, infinity_a=lay.infinity_a, infinity_b=lay.infinity_b, nb_wires=lay.nb_wires,
                      nb_public=nb_public, commitment_keys=cks, curve=curve, **lay.points, **arrs)


def read_dump_to_device(path, curve="bn254", nb_public=0, cfg=None, k_index=None):
    """A ProvingKey whose point slices went straight from the dump to the GPU (this shard's ranges only): the host
    side holds the header fields and empty point arrays.  Commitment keys are small and are read normally."""
    cfg = cfg or ProverConfig()
    g1w, g2w = _lib.CURVES[curve][1], _lib.CURVES[curve][2]
    with open(path, "rb") as f:
        lay = read_dump_layout(f, curve)
        cks = [CommitmentKey(_read_span(f, b, g1w), _read_span(f, e, g1w)) for b, e in lay.commitment_spans]
        ctx = _lib.Context(cfg.device, curve=curve)
        try:
            handle = ctx.pk_upload_from_file(f.fileno(), lay.domain_n, lay.spans, lay.infinity_a, lay.infinity_b,
                                             lay.nb_wires, k_index=k_index, shard_index=cfg.shard_index,
                                             shard_count=cfg.shard_count)
        except Exception:
            ctx.close()
            raise
    e1, e2 = np.zeros((0, g1w), dtype=np.uint64), np.zeros((0, g2w), dtype=np.uint64)
    pk = ProvingKey(domain_n=lay.domain_n, g1_a=e1, g1_b=e1, g1_z=e1, g1_k=e1, g2_b=e2, infinity_a=lay.infinity_a,
                    infinity_b=lay.infinity_b, nb_wires=lay.nb_wires, nb_public=nb_public, commitment_keys=cks,
                    curve=curve, k_index=k_index, **lay.points)
    pk.device = (ctx, handle)
    return pk


def dump_bytes(pk):
    b = io.BytesIO()
    write_dump(pk, b)
    return b.getvalue()
