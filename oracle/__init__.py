"""CPU oracle for the Groth16 proving hot path (BN254; BLS12-377 through the alias package oracle.bls12_377 and
the second build of oracle/c).

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import or execute it, and only as the checker or
the timed CPU baseline.  The accelerated path never routes through it.

Parity status: **unpinned by golden vectors** -- the reference
(/root/reference, a gnark v0.11 fork) holds no MSM / FFT / proof-byte vectors
for this path and cannot be compiled here (no Go toolchain; the arithmetic
lives in github.com/consensys/gnark-crypto
v0.14.1-0.20241010154951-6638408a49f3, not vendored).  The oracle restates the
published algorithms that the reference's call sites use
(backend/groth16/bn254/prove.go, setup.go, verify.go) and is pinned by
  * the reference's own KATs that exist for the path (filterHeap,
    backend/groth16/bn254/utils_test.go:17-38; cubic X=3,Y=35 valid / 42,42
    invalid, examples/cubic/cubic_test.go:23-38),
  * algebraic self-validation (tests/test_oracle_*.py): NTT vs O(n^2)
    evaluation, h*(X^n-1) == A*B-C, Pippenger vs double-and-add vs the trapdoor
    identity, proofs accepted by the restated pairing check of verify.go.
"""
