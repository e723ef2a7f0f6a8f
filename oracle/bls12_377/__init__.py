"""BLS12-377 instance of the oracle (BASELINE.json config 5): an alias package whose __path__ is the parent
directory, so that ``oracle.bls12_377.curve / ntt / msm / r1cs / groth16`` are second instances of the same
modules with curve.NAME == "bls12_377".  Test infrastructure only (see oracle/__init__.py).  The pairing check
(pairing.py) is BN254-only: BLS12-377 proofs are checked on the exponent side with the known toxic waste."""
import os

__path__ = [os.path.dirname(os.path.dirname(os.path.abspath(__file__)))]
