"""witness.Witness binary protocol (backend/witness/witness.go:15-38,162-240):

    [uint32(nbPublic) | uint32(nbSecret) | uint32(len(vector)) | elements]     all big-endian, elements canonical

public variables first, then secret ones, each in circuit-definition order; the ONE wire is not part of a witness.
The package comment of witness.go carries a worked example (Y = 35 public, X = 3, Z = 2 secret), which the tests use
as the golden vector.
"""
import struct
from dataclasses import dataclass

import numpy as np

from .bsb22 import FR_MOD, fr_vec_to_limbs


@dataclass
class Witness:
    nb_public: int
    nb_secret: int
    values: list            # canonical ints, public first
    curve: str = "bn254"

    def write_to(self):
        if len(self.values) != self.nb_public + self.nb_secret:
            raise ValueError("invalid witness: nbPublic + nbSecret != len(vector)")
        r = FR_MOD[self.curve]
        head = struct.pack(">III", self.nb_public, self.nb_secret, len(self.values))
        return head + b"".join((int(v) % r).to_bytes(32, "big") for v in self.values)

    @classmethod
    def read_from(cls, buf, curve="bn254"):
        """(witness, bytes read); ReadFrom's checks: the vector length must be nbPublic + nbSecret, elements < r."""
        if len(buf) < 12:
            raise ValueError("unexpected EOF")
        nb_public, nb_secret, n = struct.unpack(">III", buf[:12])
        if n != nb_public + nb_secret:
            raise ValueError("invalid witness size")
        if len(buf) < 12 + 32 * n:
            raise ValueError("unexpected EOF")
        vals = [int.from_bytes(buf[12 + 32 * i:44 + 32 * i], "big") for i in range(n)]
        if any(v >= FR_MOD[curve] for v in vals):
            raise ValueError("invalid fr.Element encoding")
        return cls(nb_public, nb_secret, vals, curve), 12 + 32 * n

    def public(self):
        """Witness.Public(): the public part as a witness of its own (what Verify takes)."""
        return Witness(self.nb_public, 0, list(self.values[:self.nb_public]), self.curve)

    def limbs(self):
        """the values as fr.Element memory images (Montgomery limbs), e.g. to seed a solver's wire vector."""
        return fr_vec_to_limbs(self.values, self.curve)
