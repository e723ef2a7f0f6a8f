"""gnark_icicle_b200 -- B200-native Groth16 (BN254) proving hot path behind gnark's icicle surface.

  csrc/        hand-written sm_100a CUDA (field core, NTT/computeH, Pippenger MSM G1/G2) + the C ABI
  lib.py       ctypes binding of libg16b200.so (stand-in for the cgo layer)
  groth16.py   mirror of package icicle_bn254 (Prove, ProvingKey, WithIcicleAcceleration)
  bsb22.py     BSB22 commitments on that path (hint override, commitment keys on the device, folded PoK)
  dump.py      ProvingKey.WriteDump / ReadDump and the direct file -> device key load
  marshal.py   Proof / VerifyingKey / ProvingKey WriteTo, WriteRawTo, ReadFrom (compressed and raw encodings, BN254)
  witness.py   witness.Witness binary protocol
  kzg.py       kzg.Commit (PLONK prover) and the MPC-setup linear combinations on the same MSM kernels

The directory ``gnark-icicle_b200`` at the repo root is a link to this package (a hyphen is not importable).
"""
from . import lib, bsb22
from .groth16 import (HasIcicle, Prove, ProvingKey, Proof, R1CSSolution, WithIcicleAcceleration, WithRandomness,
                      WithDevice, WithDevices, NewProverConfig, prove_msms, assemble, combine_shards, ConstraintSystem,
                      prove_msms_wires)
from . import dump, marshal, witness, kzg

__all__ = ["lib", "bsb22", "dump", "marshal", "witness", "kzg", "HasIcicle", "Prove", "ProvingKey", "Proof", "R1CSSolution", "WithIcicleAcceleration",
           "WithRandomness", "WithDevice", "WithDevices", "NewProverConfig", "prove_msms", "assemble", "combine_shards",
           "ConstraintSystem", "prove_msms_wires"]
