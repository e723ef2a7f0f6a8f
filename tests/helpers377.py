"""BLS12-377 test helpers: Python-oracle values <-> gnark memory image as numpy limb arrays (Fr (n,4), G1 (n,12),
G2 (n,24) uint64).  These serve the Python-oracle tests; the large-size tests (test_gpu_bls12_377_large.py) use oracle.bls12_377.cpu, the C oracle built for this curve."""
import numpy as np

from oracle.bls12_377 import curve as C

R = C.R


def fr_arr(vals):
    return np.frombuffer(C.fr_vec_to_bytes(vals), dtype=np.uint64).reshape(-1, 4).copy()


def fr_ints(arr):
    return C.fr_vec_from_bytes(np.ascontiguousarray(arr, dtype=np.uint64).tobytes())


def g1_arr(pts):
    if not len(pts):
        return np.zeros((0, 12), dtype=np.uint64)
    return np.frombuffer(C.g1_vec_to_bytes(pts), dtype=np.uint64).reshape(-1, 12).copy()


def g2_arr(pts):
    if not len(pts):
        return np.zeros((0, 24), dtype=np.uint64)
    return np.frombuffer(C.g2_vec_to_bytes(pts), dtype=np.uint64).reshape(-1, 24).copy()


def g1_pts(arr):
    b = np.ascontiguousarray(arr, dtype=np.uint64).tobytes()
    return [C.g1_from_bytes(b[i:i + 96]) for i in range(0, len(b), 96)]


def g2_pts(arr):
    b = np.ascontiguousarray(arr, dtype=np.uint64).tobytes()
    return [C.g2_from_bytes(b[i:i + 192]) for i in range(0, len(b), 192)]


def fr_rand(rng, n):
    return [rng.field(R) for _ in range(n)]


def dot(a, b):
    return sum(x * y for x, y in zip(a, b)) % R
