"""BLS12-377 instance of the oracle (BASELINE.json config 5): an alias package whose __path__ is the parent
directory, so that ``oracle.bls12_377.curve / ntt / msm / r1cs / groth16`` are second instances of the same
modules with curve.NAME == "bls12_377" -- cpu (the C oracle built with ORC_CURVE=377) and pairing (ate pairing over
Fp12 = Fp[w]/(w^12 + 5)) included.  Test infrastructure only (see oracle/__init__.py)."""
import os

__path__ = [os.path.dirname(os.path.dirname(os.path.abspath(__file__)))]
