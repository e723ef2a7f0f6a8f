"""Shared test helpers: seeded inputs and oracle-side expected values (oracle/ is the checker only)."""
import numpy as np

from oracle import cpu as ocpu
from oracle import bn254 as C
from oracle.prng import SEED_BASE

R = C.R


def fr_rand(seed, n):
    return ocpu.fr_random(SEED_BASE + seed, n)


def g1_bases(seed, n):
    """distinct points with known discrete logs: P_i = [k_i]G."""
    k = fr_rand(seed, n)
    return k, ocpu.g1_mul_gen_batch(k)


def g2_bases(seed, n):
    k = fr_rand(seed, n)
    return k, ocpu.g2_mul_gen_batch(k)


def g1_gen_arr():
    return ocpu.g1_arr([C.G1_GEN])[0]


def g2_gen_arr():
    return ocpu.g2_arr([C.G2_GEN])[0]


def trapdoor_g1(scalars, logs):
    """[sum s_i k_i] G via the oracle's double-and-add (independent of any bucket method)."""
    return ocpu.g1_mul(g1_gen_arr(), ocpu.fr_dot(scalars, logs))


def trapdoor_g2(scalars, logs):
    return ocpu.g2_mul(g2_gen_arr(), ocpu.fr_dot(scalars, logs))


def fr_from_ints(vals):
    return ocpu.fr_arr([v % R for v in vals])


def assert_limbs_equal(got, want, what=""):
    got = np.asarray(got, dtype=np.uint64).reshape(-1)
    want = np.asarray(want, dtype=np.uint64).reshape(-1)
    assert got.shape == want.shape, (what, got.shape, want.shape)
    if not np.array_equal(got, want):
        bad = np.nonzero(got != want)[0]
        raise AssertionError(f"{what}: {len(bad)} of {got.size} limbs differ, first at limb {bad[0]} "
                             f"(element {bad[0] // 4}): got {got[bad[0]]:#x} want {want[bad[0]]:#x}")


# ---------------------------------------------------------------- proving keys for the prove tests
def bitrev_perm(n):
    logn = n.bit_length() - 1
    idx = np.arange(n, dtype=np.uint64)
    rev = np.zeros(n, dtype=np.uint64)
    for b in range(logn):
        rev |= ((idx >> np.uint64(b)) & np.uint64(1)) << np.uint64(logn - 1 - b)
    return rev.astype(np.int64)


def pk_from_python(pk_py, nb_wires):
    """oracle.groth16.setup() result -> gnark_icicle_b200.ProvingKey (numpy limb arrays)."""
    from gnark_icicle_b200 import ProvingKey
    g1 = lambda pts: ocpu.g1_arr(pts) if len(pts) else np.zeros((0, 8), dtype=np.uint64)
    g2 = lambda pts: ocpu.g2_arr(pts) if len(pts) else np.zeros((0, 16), dtype=np.uint64)
    return ProvingKey(
        domain_n=pk_py.domain.n,
        g1_alpha=g1([pk_py.g1_alpha])[0], g1_beta=g1([pk_py.g1_beta])[0], g1_delta=g1([pk_py.g1_delta])[0],
        g1_a=g1(pk_py.g1_a), g1_b=g1(pk_py.g1_b), g1_z=g1(pk_py.g1_z), g1_k=g1(pk_py.g1_k),
        g2_beta=g2([pk_py.g2_beta])[0], g2_delta=g2([pk_py.g2_delta])[0], g2_b=g2(pk_py.g2_b),
        infinity_a=np.array(pk_py.infinity_a, dtype=np.uint8), infinity_b=np.array(pk_py.infinity_b, dtype=np.uint8),
        nb_wires=nb_wires, nb_public=pk_py.nb_public)


def is_zero_rows(a):
    return ~np.any(a != 0, axis=1)


class ChainCase:
    """MiMC-shaped chain (config 4 shape) with a trapdoor setup: scalars from the C oracle, bases = [scalar]G
    produced by `g1_mul` / `g2_mul` callables (CPU oracle for small sizes, the GPU fixed-base kernel for large)."""

    def __init__(self, nc, seed, g1_mul=None, g2_mul=None):
        from gnark_icicle_b200 import ProvingKey
        g1_mul = g1_mul or ocpu.g1_mul_gen_batch
        g2_mul = g2_mul or ocpu.g2_mul_gen_batch
        n = 1
        while n < nc:
            n <<= 1
        self.nc, self.n, self.nw = nc, n, nc + 3
        rnd = fr_rand(1000 + seed, 9)
        self.tw = rnd[0:5]
        self.r, self.s = rnd[5], rnd[6]
        x0, k = rnd[7], rnd[8]
        cseed = SEED_BASE + 2000 + seed
        self.W, self.a, self.b, self.c = ocpu.mimc_chain_witness(nc, cseed, x0, k)
        A, B, K, Z = ocpu.mimc_chain_setup_scalars(nc, n, cseed, self.tw)
        self.inf_a, self.inf_b = is_zero_rows(A), is_zero_rows(B)
        self.sA, self.sB = A[~self.inf_a], B[~self.inf_b]
        self.sK = K[2:]
        self.sZ = Z[bitrev_perm(n)][: n - 1]          # setup.go:265-267
        abd = np.stack([self.tw[1], self.tw[2], self.tw[4]])
        g1 = g1_mul(np.concatenate([abd, self.sA, self.sB, self.sK, self.sZ]))
        o = 3
        pa = g1[o:o + len(self.sA)]; o += len(self.sA)
        pb = g1[o:o + len(self.sB)]; o += len(self.sB)
        pkk = g1[o:o + len(self.sK)]; o += len(self.sK)
        pz = g1[o:o + len(self.sZ)]; o += len(self.sZ)
        g2 = g2_mul(np.concatenate([np.stack([self.tw[2], self.tw[4]]), self.sB]))
        self.pk = ProvingKey(domain_n=n, g1_alpha=g1[0], g1_beta=g1[1], g1_delta=g1[2], g1_a=pa, g1_b=pb, g1_z=pz,
                             g1_k=pkk, g2_beta=g2[0], g2_delta=g2[1], g2_b=g2[2:], infinity_a=self.inf_a.astype(np.uint8),
                             infinity_b=self.inf_b.astype(np.uint8), nb_wires=self.nw, nb_public=2)

    def expected_points(self, h):
        """the five raw MSM results through the trapdoor identity (exponent side), given the oracle's h."""
        wa, wb, wk = self.W[~self.inf_a], self.W[~self.inf_b], self.W[2:]
        return np.concatenate([
            trapdoor_g1(wa, self.sA), trapdoor_g1(wb, self.sB), trapdoor_g1(wk, self.sK),
            trapdoor_g1(h[: self.n - 1], self.sZ), trapdoor_g2(wb, self.sB)])
