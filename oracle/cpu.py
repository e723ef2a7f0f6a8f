"""ctypes binding of the C oracle (oracle/c/liboracle_cpu*.so).  Test infrastructure only.

Arrays are numpy uint64 with a trailing limb axis: Fr (n,4), G1 affine (n,G1W),
G2 affine (n,G2W) -- the gnark memory image (Montgomery, little-endian limbs); G1W, G2W = 8, 16 for BN254 and
12, 24 for BLS12-377.  Curve-generic like oracle/curve.py: ``oracle.cpu`` binds the BN254 library and
``oracle.bls12_377.cpu`` (the alias package) the BLS12-377 one, both built from the same C sources.
"""
import ctypes
import os
import subprocess
import numpy as np

NAME = "bls12_377" if (__package__ or "").endswith("bls12_377") else "bn254"
G1W, G2W = (12, 24) if NAME == "bls12_377" else (8, 16)
_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "c")
_SO_NAME = "liboracle_cpu_bls12_377.so" if NAME == "bls12_377" else "liboracle_cpu.so"
_SO = os.path.join(_DIR, _SO_NAME)


def build(force=False):
    srcs = [os.path.join(_DIR, f) for f in ("oracle_cpu.c", "field.h", "field_tmpl.h", "ec_tmpl.h", "Makefile")]
    if force or not os.path.exists(_SO) or any(
            os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _DIR, "-B", _SO_NAME], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.orc_threads.restype = ctypes.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _c(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


u64 = ctypes.c_uint64
ci = ctypes.c_int


def threads():
    return lib().orc_threads()


def fr_random(seed, n):
    out = np.empty((n, 4), dtype=np.uint64)
    lib().orc_fr_random(u64(seed), u64(n), _p(out))
    return out


def fr_dot(a, b):
    a, b = _c(a), _c(b)
    out = np.empty(4, dtype=np.uint64)
    lib().orc_fr_dot(_p(a), _p(b), u64(len(a)), _p(out))
    return out


def fr_mul_vec(a, b):
    a, b = _c(a), _c(b)
    out = np.empty_like(a)
    lib().orc_fr_mul_vec(_p(a), _p(b), u64(len(a)), _p(out))
    return out


def g1_mul_gen_batch(scalars, nthreads=0):
    s = _c(scalars)
    out = np.empty((len(s), G1W), dtype=np.uint64)
    lib().orc_g1_mul_gen_batch(_p(s), u64(len(s)), _p(out), ci(nthreads or threads()))
    return out


def g2_mul_gen_batch(scalars, nthreads=0):
    s = _c(scalars)
    out = np.empty((len(s), G2W), dtype=np.uint64)
    lib().orc_g2_mul_gen_batch(_p(s), u64(len(s)), _p(out), ci(nthreads or threads()))
    return out


def msm_g1(points, scalars, c=0, nthreads=0):
    p, s = _c(points), _c(scalars)
    assert len(p) == len(s)
    out = np.zeros(G1W, dtype=np.uint64)
    lib().orc_msm_g1(_p(p), _p(s), u64(len(s)), _p(out), ci(c), ci(nthreads or threads()))
    return out


def msm_g2(points, scalars, c=0, nthreads=0):
    p, s = _c(points), _c(scalars)
    assert len(p) == len(s)
    out = np.zeros(G2W, dtype=np.uint64)
    lib().orc_msm_g2(_p(p), _p(s), u64(len(s)), _p(out), ci(c), ci(nthreads or threads()))
    return out


def g1_mul(point, k):
    p, k = _c(point), _c(k)
    out = np.zeros(G1W, dtype=np.uint64)
    lib().orc_g1_mul(_p(p), _p(k), _p(out))
    return out


def g2_mul(point, k):
    p, k = _c(point), _c(k)
    out = np.zeros(G2W, dtype=np.uint64)
    lib().orc_g2_mul(_p(p), _p(k), _p(out))
    return out


def g1_add(a, b):
    a, b = _c(a), _c(b)
    out = np.zeros(G1W, dtype=np.uint64)
    lib().orc_g1_add(_p(a), _p(b), _p(out))
    return out


def g2_add(a, b):
    a, b = _c(a), _c(b)
    out = np.zeros(G2W, dtype=np.uint64)
    lib().orc_g2_add(_p(a), _p(b), _p(out))
    return out


FFT_INV_DIF, FFT_DIT_COSET, FFT_INV_DIF_COSET, FFT_DIT, FFT_DIF = 0, 1, 2, 3, 4


def fft(a, mode, nthreads=0):
    a = _c(a).copy()
    lib().orc_fft(_p(a), u64(len(a)), ci(mode), ci(nthreads or threads()))
    return a


def compute_h(a, b, c, n, nthreads=0):
    a, b, c = _c(a), _c(b), _c(c)
    h = np.empty((n, 4), dtype=np.uint64)
    lib().orc_compute_h(_p(a), _p(b), _p(c), u64(len(a)), u64(n), _p(h), ci(nthreads or threads()))
    return h


def mimc_chain_witness(nc, seed, x0, k):
    x0, k = _c(x0), _c(k)
    W = np.empty((nc + 3, 4), dtype=np.uint64)
    a, b, c = (np.empty((nc, 4), dtype=np.uint64) for _ in range(3))
    lib().orc_mimc_chain_witness(u64(nc), u64(seed), _p(x0), _p(k), _p(W), _p(a), _p(b), _p(c))
    return W, a, b, c


def mimc_chain_constants(nc, seed):
    lib().orc_mimc_chain_per_round.restype = ctypes.c_int
    out = np.empty(((nc - 1) // lib().orc_mimc_chain_per_round(), 4), dtype=np.uint64)
    lib().orc_mimc_chain_constants(u64(nc), u64(seed), _p(out))
    return out


def mimc_chain_setup_scalars(nc, n, seed, tw):
    """tw: (5,4) Montgomery {t, alpha, beta, gamma, delta}. Returns A,B,K (nc+3), Z (n, natural order)."""
    tw = _c(tw)
    A, B, K = (np.empty((nc + 3, 4), dtype=np.uint64) for _ in range(3))
    Z = np.empty((n, 4), dtype=np.uint64)
    lib().orc_mimc_chain_setup_scalars(u64(nc), u64(n), u64(seed), _p(tw), _p(A), _p(B), _p(K), _p(Z), ci(threads()))
    return A, B, K, Z


# ---- conversions between Python ints (oracle/curve.py of the same instance) and limb arrays
def fr_arr(vals):
    from .curve import fr_vec_to_bytes
    return np.frombuffer(fr_vec_to_bytes(vals), dtype=np.uint64).reshape(-1, 4).copy()


def fr_ints(arr):
    from .curve import fr_vec_from_bytes
    return fr_vec_from_bytes(_c(arr).tobytes())


def g1_arr(pts):
    from .curve import g1_vec_to_bytes
    return np.frombuffer(g1_vec_to_bytes(pts), dtype=np.uint64).reshape(-1, G1W).copy()


def g2_arr(pts):
    from .curve import g2_vec_to_bytes
    return np.frombuffer(g2_vec_to_bytes(pts), dtype=np.uint64).reshape(-1, G2W).copy()


def g1_pt(arr):
    from .curve import g1_from_bytes
    return g1_from_bytes(_c(arr).tobytes())


def g2_pt(arr):
    from .curve import g2_from_bytes
    return g2_from_bytes(_c(arr).tobytes())


def g1_gen_arr():
    from .curve import G1_GEN
    return g1_arr([G1_GEN])[0]


def g2_gen_arr():
    from .curve import G2_GEN
    return g2_arr([G2_GEN])[0]
