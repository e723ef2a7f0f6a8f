"""Other consumers of the MSM kernels: KZG commitments of the PLONK prover and the linear
combinations of the Groth16 MPC setup.  Host-side mirrors above the same C ABI the Groth16 path uses.

  kzg.Commit(p, pk)                 backend/plonk/bn254/prove.go:312,460,718 (Bsb22 / wire / Z commitments on
                                    pk.KzgLagrange or pk.Kzg), :1158-1172 commitToQuotient -- gnark-crypto's
                                    kzg.Commit is MultiExp(pk.G1[:len(p)], p) and fails with ErrInvalidPolynomialSize
                                    for an empty polynomial or one longer than the SRS  [gnark-crypto: recollection]
  merge / linearCombinationG1/G2    backend/groth16/bn254/mpcsetup/utils.go:120-154 (sum r_i A_i, sum r_i A_{i+1};
                                    the reference samples r itself, here it is an argument so results can be compared)

The SRS of a KZG key is fixed across proofs, so it stays resident on the device with its table of multiples
(g16_bases_upload) exactly like the Groth16 key; the MPC-setup bases change every call and go through g16_msm_g1/g2.
"""
from dataclasses import dataclass

import numpy as np

from . import lib as _lib


class ErrInvalidPolynomialSize(ValueError):
    """kzg.ErrInvalidPolynomialSize: "invalid polynomial size (larger than SRS or == 0)"."""


@dataclass
class ProvingKey:
    """kzg.ProvingKey: G1 = [G, [tau]G, [tau^2]G, ...] (or the Lagrange-basis SRS), gnark memory image (n, 8)."""
    G1: np.ndarray
    device: object = None      # (Context, bases handle)

    def setup_device(self, ctx=None):
        if self.device is not None and (ctx is None or self.device[0] is ctx):
            return self.device
        self.free_device()
        ctx = ctx or _lib.Context(0)
        self.device = (ctx, ctx.bases_upload(self.G1))
        return self.device

    def free_device(self):
        if self.device is not None:
            ctx, h = self.device
            ctx.bases_free(h)
            self.device = None


def Commit(p, pk, ctx=None):
    """kzg.Commit: digest = sum_i p_i * pk.G1[i] as a G1Affine memory image (8 limbs)."""
    p = np.ascontiguousarray(p, dtype=np.uint64).reshape(-1, 4)
    n = len(pk.G1)
    if len(p) == 0 or len(p) > n:
        raise ErrInvalidPolynomialSize("invalid polynomial size (larger than SRS or == 0)")
    c, h = pk.setup_device(ctx)
    if len(p) < n:                         # MultiExp over pk.G1[:len(p)]: the missing scalars are zero
        p = np.concatenate([p, np.zeros((n - len(p), 4), dtype=np.uint64)])
    return c.msm_resident(h, p)


def commit_to_quotient(h1, h2, h3, pk, ctx=None):
    """commitToQuotient (prove.go:1158-1172): the three quotient shards against the same key."""
    return [Commit(h, pk, ctx) for h in (h1, h2, h3)]


def merge(A, B, r, ctx):
    """mpcsetup/utils.go:120-130: a = sum r_i A_i, b = sum r_i B_i."""
    return ctx.msm_g1(r, A), ctx.msm_g1(r, B)


def linear_combination_g1(A, r, ctx):
    """mpcsetup/utils.go:132-142: L1 = sum r_i A_i, L2 = sum r_i A_{i+1}, r of length len(A) - 1."""
    A = np.ascontiguousarray(A, dtype=np.uint64)
    return ctx.msm_g1(r, A[:-1]), ctx.msm_g1(r, A[1:])


def linear_combination_g2(A, r, ctx):
    """mpcsetup/utils.go:144-154, over G2."""
    A = np.ascontiguousarray(A, dtype=np.uint64)
    return ctx.msm_g2(r, A[:-1]), ctx.msm_g2(r, A[1:])
