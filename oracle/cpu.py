"""ctypes binding of the C oracle (oracle/c/liboracle_cpu.so).  Test infrastructure only.

Arrays are numpy uint64 with a trailing limb axis: Fr (n,4), G1 affine (n,8),
G2 affine (n,16) -- the gnark memory image (Montgomery, little-endian limbs).
"""
import ctypes
import os
import subprocess
import numpy as np

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "c")
_SO = os.path.join(_DIR, "liboracle_cpu.so")


def build(force=False):
    srcs = [os.path.join(_DIR, f) for f in ("oracle_cpu.c", "field.h", "ec_tmpl.h", "Makefile")]
    if force or not os.path.exists(_SO) or any(
            os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _DIR, "-B", "liboracle_cpu.so"], stdout=subprocess.DEVNULL)
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
    out = np.empty((len(s), 8), dtype=np.uint64)
    lib().orc_g1_mul_gen_batch(_p(s), u64(len(s)), _p(out), ci(nthreads or threads()))
    return out


def g2_mul_gen_batch(scalars, nthreads=0):
    s = _c(scalars)
    out = np.empty((len(s), 16), dtype=np.uint64)
    lib().orc_g2_mul_gen_batch(_p(s), u64(len(s)), _p(out), ci(nthreads or threads()))
    return out


def msm_g1(points, scalars, c=0, nthreads=0):
    p, s = _c(points), _c(scalars)
    assert len(p) == len(s)
    out = np.zeros(8, dtype=np.uint64)
    lib().orc_msm_g1(_p(p), _p(s), u64(len(s)), _p(out), ci(c), ci(nthreads or threads()))
    return out


def msm_g2(points, scalars, c=0, nthreads=0):
    p, s = _c(points), _c(scalars)
    assert len(p) == len(s)
    out = np.zeros(16, dtype=np.uint64)
    lib().orc_msm_g2(_p(p), _p(s), u64(len(s)), _p(out), ci(c), ci(nthreads or threads()))
    return out


def g1_mul(point, k):
    p, k = _c(point), _c(k)
    out = np.zeros(8, dtype=np.uint64)
    lib().orc_g1_mul(_p(p), _p(k), _p(out))
    return out


def g2_mul(point, k):
    p, k = _c(point), _c(k)
    out = np.zeros(16, dtype=np.uint64)
    lib().orc_g2_mul(_p(p), _p(k), _p(out))
    return out


def g1_add(a, b):
    a, b = _c(a), _c(b)
    out = np.zeros(8, dtype=np.uint64)
    lib().orc_g1_add(_p(a), _p(b), _p(out))
    return out


def g2_add(a, b):
    a, b = _c(a), _c(b)
    out = np.zeros(16, dtype=np.uint64)
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
    out = np.empty(((nc - 1) // 3, 4), dtype=np.uint64)
    lib().orc_mimc_chain_constants(u64(nc), u64(seed), _p(out))
    return out


def mimc_chain_setup_scalars(nc, n, seed, tw):
    """tw: (5,4) Montgomery {t, alpha, beta, gamma, delta}. Returns A,B,K (nc+3), Z (n, natural order)."""
    tw = _c(tw)
    A, B, K = (np.empty((nc + 3, 4), dtype=np.uint64) for _ in range(3))
    Z = np.empty((n, 4), dtype=np.uint64)
    lib().orc_mimc_chain_setup_scalars(u64(nc), u64(n), u64(seed), _p(tw), _p(A), _p(B), _p(K), _p(Z), ci(threads()))
    return A, B, K, Z


# ---- conversions between Python ints (oracle/bn254.py) and limb arrays
def fr_arr(vals):
    from .bn254 import fr_vec_to_bytes
    return np.frombuffer(fr_vec_to_bytes(vals), dtype=np.uint64).reshape(-1, 4).copy()


def fr_ints(arr):
    from .bn254 import fr_vec_from_bytes
    return fr_vec_from_bytes(_c(arr).tobytes())


def g1_arr(pts):
    from .bn254 import g1_vec_to_bytes
    return np.frombuffer(g1_vec_to_bytes(pts), dtype=np.uint64).reshape(-1, 8).copy()


def g2_arr(pts):
    from .bn254 import g2_vec_to_bytes
    return np.frombuffer(g2_vec_to_bytes(pts), dtype=np.uint64).reshape(-1, 16).copy()


def g1_pt(arr):
    from .bn254 import g1_from_bytes
    return g1_from_bytes(_c(arr).tobytes())


def g2_pt(arr):
    from .bn254 import g2_from_bytes
    return g2_from_bytes(_c(arr).tobytes())
