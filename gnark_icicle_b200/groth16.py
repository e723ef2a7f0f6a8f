"""Host-side mirror of package ``icicle_bn254`` (reference: backend/groth16/bn254/icicle/{icicle.go,
provingkey.go,noicicle.go}) above the libg16b200 C ABI.

The reference's host language is Go; no Go toolchain exists on the build or GPU machines, so this mirror is
Python (the Go shim a maintainer would use is go/icicle_bn254/ + INTEGRATION.md).  Names and argument
meaning follow the reference:

  HasIcicle                                  icicle.go:29
  ProvingKey{ProvingKey; *deviceInfo}        icicle/provingkey.go:25-28   -> ProvingKey (+ .device)
  (pk *ProvingKey) setupDevicePointers()     icicle.go:31-130             -> ProvingKey.setup_device_pointers
  Prove(r1cs, pk, fullWitness, opts...)      icicle.go:133-422            -> Prove
  backend.WithIcicleAcceleration()           backend/backend.go:142-147   -> WithIcicleAcceleration

Differences that are deliberate: r, s can be injected (the reference hard-codes SetRandom at
icicle.go:283-288, which makes byte-exact comparison impossible); the BSB22 commitment MSMs (prove.go:94,124,
which the reference leaves on the CPU) run on the device against resident commitment keys (bsb22.py).
"""
import secrets
from dataclasses import dataclass, field as _field

import numpy as np

from . import lib as _lib
from . import bsb22 as _bsb22

HasIcicle = True
_FR_MOD = {"bn254": 21888242871839275222246405745257275088548364400416034343698204186575808495617,
           "bls12_377": 8444461749428370424248824938781546531375899335154063827935233455917409239041}


# ---- backend.ProverOption (backend/backend.go:66-147)
@dataclass
class ProverConfig:
    accelerator: str = ""
    r: object = None          # injected randomness (Fr limbs, Montgomery); None = sample
    s: object = None
    device: int = 0
    shard_index: int = 0
    shard_count: int = 1


def WithIcicleAcceleration():
    def opt(cfg):
        cfg.accelerator = "icicle"
    return opt


def WithRandomness(r, s):
    def opt(cfg):
        cfg.r, cfg.s = r, s
    return opt


def WithDevice(device, shard_index=0, shard_count=1):
    def opt(cfg):
        cfg.device, cfg.shard_index, cfg.shard_count = device, shard_index, shard_count
    return opt


def WithDevices(devices):
    """2, 4 or 8 GPU ordinals: one multi-GPU context behind the same Prove call (g16_ctx_create over several devices)."""
    def opt(cfg):
        cfg.device, cfg.shard_index, cfg.shard_count = list(devices), 0, 1
    return opt


def NewProverConfig(*opts):
    cfg = ProverConfig()
    for o in opts:
        o(cfg)
    return cfg


# ---- data crossing the boundary
@dataclass
class R1CSSolution:
    """constraint/bn254/system.go:221-224: W (all wires) and A, B, C (L.w, R.w, O.w per constraint)."""
    W: np.ndarray
    A: np.ndarray
    B: np.ndarray
    C: np.ndarray


@dataclass
class ConstraintSystem:
    """The part of cs.R1CS (constraint/bn254/system.go, constraint/r1cs.go) the device needs to produce a = L.w,
    b = R.w, c = O.w itself: the coefficient table and the three matrices in CSR form over the
    constraints, a term being (coefficient id, wire id) like constraint.Term{CID, VID}.  Uploaded once, lazily,
    next to the proving key; a Prove call then only ships the wire vector."""
    nb_constraints: int
    nb_wires: int
    coefficients: np.ndarray          # (k, 4) Fr, Montgomery
    L: tuple                          # (row_ptr, wire, coeff_id) uint32 arrays
    R: tuple
    O: tuple
    device: object = None             # (Context, handle)

    def setup_device(self, ctx):
        if self.device is not None and self.device[0] is ctx:
            return self.device[1]
        self.free_device()
        h = ctx.r1cs_upload(self.nb_constraints, self.nb_wires, self.coefficients, (self.L, self.R, self.O))
        self.device = (ctx, h)
        return h

    def free_device(self):
        if self.device is not None:
            ctx, h = self.device
            ctx.r1cs_free(h)
            self.device = None


@dataclass
class Proof:
    """prove.go:44-49: raw gnark memory images."""
    Ar: np.ndarray    # (8,)  G1Affine   ((12,) for BLS12-377)
    Bs: np.ndarray    # (16,) G2Affine   ((24,))
    Krs: np.ndarray   # (8,)  G1Affine
    Commitments: list = _field(default_factory=list)    # []G1Affine, one per BSB22 commitment
    CommitmentPok: np.ndarray = None                    # G1Affine, the folded proof of knowledge

    def raw(self):
        """Ar || Bs || Krs || Commitments... || CommitmentPok (the last two only when there are commitments)."""
        parts = [self.Ar, self.Bs, self.Krs] + list(self.Commitments)
        if self.Commitments:
            parts.append(self.CommitmentPok)
        return np.concatenate(parts).tobytes()

    def write_to(self):
        """Proof.WriteTo (marshal.go:30-35): compressed points."""
        from . import marshal
        return marshal.proof_write_to(self, raw=False)

    def write_raw_to(self):
        """Proof.WriteRawTo (marshal.go:37-42)."""
        from . import marshal
        return marshal.proof_write_to(self, raw=True)


@dataclass
class ProvingKey:
    """setup.go:35-58 fields that feed the prover, as numpy limb arrays, plus the device info."""
    domain_n: int
    g1_alpha: np.ndarray
    g1_beta: np.ndarray
    g1_delta: np.ndarray
    g1_a: np.ndarray
    g1_b: np.ndarray
    g1_z: np.ndarray
    g1_k: np.ndarray
    g2_beta: np.ndarray
    g2_delta: np.ndarray
    g2_b: np.ndarray
    infinity_a: np.ndarray
    infinity_b: np.ndarray
    nb_wires: int
    nb_public: int
    commitment_keys: list = _field(default_factory=list)   # []bsb22.CommitmentKey (pedersen.ProvingKey, setup.go:57)
    device: object = None     # deviceInfo (provingkey.go:10-23): (Context, handle)
    curve: str = "bn254"      # "bls12_377": 12/24-limb points, the g16_bls12_377_* entry points (config 5)
    k_index: np.ndarray = None           # wire feeding each K base; None = W[nb_wires - len(g1_k):] (no commitments)
    commitment_device: object = None     # (Context, [commitment-key handles])

    def setup_device_pointers(self, cfg=None):
        """icicle.go:31-130: one-time upload; lazily on first Prove (icicle.go:145-150)."""
        if self.device is not None:
            return
        cfg = cfg or ProverConfig()
        ctx = _lib.Context(cfg.device, curve=self.curve)
        handle = ctx.pk_upload(self.domain_n, self.g1_a, self.g1_b, self.g1_k, self.g1_z, self.g2_b,
                               self.infinity_a, self.infinity_b, self.nb_wires, k_index=self.k_index,
                               shard_index=cfg.shard_index, shard_count=cfg.shard_count)
        self.device = (ctx, handle)

    def pk_points(self):
        return np.concatenate([self.g1_alpha.reshape(-1), self.g1_beta.reshape(-1), self.g1_delta.reshape(-1),
                               self.g2_beta.reshape(-1), self.g2_delta.reshape(-1)]).astype(np.uint64)

    def free_device(self):
        if self.device is not None:
            ctx, handle = self.device
            if self.commitment_device is not None:
                for h in self.commitment_device[1]:
                    ctx.commitment_key_free(h)
                self.commitment_device = None
            ctx.pk_free(handle)
            ctx.close()
            self.device = None


def _random_fr(curve="bn254"):
    v = secrets.randbelow(_FR_MOD[curve])
    return np.array([(v >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)], dtype=np.uint64)


def prove_msms(pk, sol, cfg=None):
    """Device section of Prove (icicle.go:204-410): returns (48 limbs = 5 raw MSM points, stats)."""
    pk.setup_device_pointers(cfg)
    ctx, handle = pk.device
    return ctx.prove(handle, sol.W, sol.A, sol.B, sol.C)


def prove_msms_wires(pk, cs, W, cfg=None):
    """prove_msms with a, b, c evaluated on the device from the resident constraint matrices."""
    pk.setup_device_pointers(cfg)
    ctx, handle = pk.device
    return ctx.prove_wires(handle, cs.setup_device(ctx), W)


def Prove(r1cs, pk, full_witness, *opts):
    """icicle_bn254.Prove (icicle.go:133-422).

    ``r1cs`` must offer ``solve(full_witness) -> R1CSSolution`` (constraint/bn254/system.go:65; the solver
    is host code outside the accelerated path) -- or pass ``r1cs=None`` and an R1CSSolution as the witness, or a
    ``ConstraintSystem`` and the solved wire vector (a, b, c are then computed on the device).
    A constraint system with BSB22 commitments additionally carries ``commitment_info`` (a list of
    bsb22.Groth16Commitment, = r1cs.CommitmentInfo) and ``nb_public``, and its ``solve(full_witness, bsb22_hint)``
    calls ``bsb22_hint(i, hashed_values, committed_values) -> int`` where gnark's solver would call the overridden
    Bsb22CommitmentComputePlaceholder hint (prove.go:81-109).
    Without WithIcicleAcceleration() the reference falls back to the CPU prover (icicle.go:141-143); this
    package has no CPU prover, so that is an error here.
    """
    cfg = NewProverConfig(*opts)
    if cfg.accelerator != "icicle":
        raise _lib.G16Error("accelerator option missing: pass WithIcicleAcceleration() "
                            "(the CPU prover lives in gnark, not in this package)")
    if cfg.shard_count > 1:
        # one shard's MSM points are partial sums: assembling them would give a proof that can never verify
        raise _lib.G16Error("Prove needs the whole key: WithDevice(dev, i, n) with n > 1 uploads one shard only. Use a "
                            "multi-GPU context (WithDevices([...])), or prove_msms per shard + combine_shards + assemble")
    info = list(getattr(r1cs, "commitment_info", None) or [])
    if info or pk.commitment_keys:
        return _prove_with_commitments(r1cs, pk, full_witness, info, cfg)
    if isinstance(r1cs, ConstraintSystem):
        # the wire vector is the whole witness-side input: the device evaluates the constraint matrices
        points, stats = prove_msms_wires(pk, r1cs, full_witness, cfg)
    else:
        sol = full_witness if r1cs is None else r1cs.solve(full_witness)
        points, stats = prove_msms(pk, sol, cfg)
    r = cfg.r if cfg.r is not None else _random_fr(pk.curve)
    s = cfg.s if cfg.s is not None else _random_fr(pk.curve)
    proof = assemble(pk, points, r, s)
    proof.stats = stats
    return proof


def _prove_with_commitments(r1cs, pk, full_witness, info, cfg):
    """Prove for a circuit that uses api.Commit: prove.go:74-139 around the usual device section."""
    if r1cs is None or isinstance(r1cs, ConstraintSystem) or not hasattr(r1cs, "solve"):
        raise _lib.G16Error("a constraint system with commitments must be solved inside Prove: the bsb22 hint "
                            "computes the commitments (prove.go:81-109); pass an r1cs with solve(witness, bsb22_hint)")
    if pk.k_index is None:
        # wires that reach the K MSM (prove.go:241-246); fixed per circuit, so it becomes part of the resident key
        if pk.device is not None:
            raise _lib.G16Error("proving key was uploaded before its commitment info was known; free_device() first")
        pk.k_index = _bsb22.k_wire_indices(r1cs.nb_public, pk.nb_wires, info)
        if len(pk.k_index) != len(pk.g1_k):
            raise _lib.G16Error(f"commitment info leaves {len(pk.k_index)} private wires, pk.G1.K has {len(pk.g1_k)}")
    pk.setup_device_pointers(cfg)
    ctx, handle = pk.device
    session = _bsb22.Session(pk, info, ctx)
    sol = r1cs.solve(full_witness, session.hint)
    points, stats = ctx.prove(handle, sol.W, sol.A, sol.B, sol.C)
    commitment_values = [_fr_from_limbs(sol.W[c.commitment_index], pk.curve) for c in info]
    commitments, pok = session.finish(commitment_values)
    r = cfg.r if cfg.r is not None else _random_fr(pk.curve)
    s = cfg.s if cfg.s is not None else _random_fr(pk.curve)
    proof = assemble(pk, points, r, s)
    proof.Commitments, proof.CommitmentPok = commitments, pok
    proof.stats = stats
    return proof


def _fr_from_limbs(limbs, curve):
    m = _FR_MOD[curve]
    v = 0
    for i, w in enumerate(np.asarray(limbs, dtype=np.uint64).reshape(-1)):
        v |= int(w) << (64 * i)
    return v * pow(1 << 256, -1, m) % m


def assemble(pk, msm_points, r, s):
    """host epilogue (icicle.go:281-295,306-307,319-321,362-372,386-391) on already-combined MSM points."""
    out = _lib.assemble_proof(pk.pk_points(), r, s, msm_points, curve=pk.curve)
    _, w1, w2 = _lib.CURVES[pk.curve]
    return Proof(Ar=out[0:w1].copy(), Bs=out[w1:w1 + w2].copy(), Krs=out[w1 + w2:].copy())


def combine_shards(point_blocks, curve="bn254"):
    """Sum the per-shard raw MSM blocks (each 48 limbs for BN254) of a point-range-sharded prove."""
    _, w1, w2 = _lib.CURVES[curve]
    acc = np.array(point_blocks[0], dtype=np.uint64).copy()
    for blk in point_blocks[1:]:
        blk = np.asarray(blk, dtype=np.uint64)
        for i in range(4):
            acc[w1 * i:w1 * i + w1] = _lib.g1_add(acc[w1 * i:w1 * i + w1], blk[w1 * i:w1 * i + w1], curve=curve)
        acc[4 * w1:] = _lib.g2_add(acc[4 * w1:], blk[4 * w1:], curve=curve)
    return acc
