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
    """MiMC-shaped chain (config 4 shape; pow17 rounds = config 5 shape with the BLS12-377 oracle) with a trapdoor
    setup: scalars from the C oracle `oc` (oracle.cpu, or oracle.bls12_377.cpu), bases = [scalar]G produced by
    `g1_mul` / `g2_mul` callables (CPU oracle for small sizes, the GPU fixed-base kernel for large)."""

    def __init__(self, nc, seed, g1_mul=None, g2_mul=None, oc=None):
        from gnark_icicle_b200 import ProvingKey
        oc = self.oc = oc or ocpu
        g1_mul = g1_mul or oc.g1_mul_gen_batch
        g2_mul = g2_mul or oc.g2_mul_gen_batch
        n = 1
        while n < nc:
            n <<= 1
        self.nc, self.n, self.nw = nc, n, nc + 3
        rnd = oc.fr_random(SEED_BASE + 1000 + seed, 9)
        self.tw = rnd[0:5]
        self.r, self.s = rnd[5], rnd[6]
        x0, k = rnd[7], rnd[8]
        cseed = SEED_BASE + 2000 + seed
        self.W, self.a, self.b, self.c = oc.mimc_chain_witness(nc, cseed, x0, k)
        A, B, K, Z = oc.mimc_chain_setup_scalars(nc, n, cseed, self.tw)
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
        oc = self.oc
        g1, g2 = oc.g1_gen_arr(), oc.g2_gen_arr()
        wa, wb, wk = self.W[~self.inf_a], self.W[~self.inf_b], self.W[2:]
        t1 = lambda sc, logs: oc.g1_mul(g1, oc.fr_dot(sc, logs))
        return np.concatenate([
            t1(wa, self.sA), t1(wb, self.sB), t1(wk, self.sK), t1(h[: self.n - 1], self.sZ),
            oc.g2_mul(g2, oc.fr_dot(wb, self.sB))])


# ---------------------------------------------------------------- constraint matrices for the device (8(f) N1)
def csr_from_oracle(cs, fr_arr):
    """oracle.r1cs.R1CS -> gnark_icicle_b200.ConstraintSystem (coefficient table de-duplicated like cs.Coefficients)."""
    from gnark_icicle_b200 import ConstraintSystem
    table, ids = [], {}
    mats = []
    for m in range(3):
        rp, wi, ci = [0], [], []
        for row in cs.constraints:
            for cf, w in row[m]:
                cf %= R
                if cf not in ids:
                    ids[cf] = len(table)
                    table.append(cf)
                wi.append(w)
                ci.append(ids[cf])
            rp.append(len(wi))
        mats.append(tuple(np.array(x, dtype=np.uint32) for x in (rp, wi, ci)))
    return ConstraintSystem(nb_constraints=cs.nb_constraints, nb_wires=cs.nb_wires, coefficients=fr_arr(table),
                            L=mats[0], R=mats[1], O=mats[2])


def chain_csr(nc, cseed):
    """the constraint matrices of ChainCase's MiMC-shaped chain (oracle/c orc_mimc_chain_*), built with numpy."""
    from gnark_icicle_b200 import ConstraintSystem
    consts = ocpu.mimc_chain_constants(nc, cseed)
    rounds = len(consts)
    one = ocpu.fr_arr([1])
    coeffs = np.concatenate([one, consts]) if rounds else one
    r = np.arange(rounds, dtype=np.int64)
    xw = np.where(r == 0, 2, 3 * r + 3)            # state wire entering round r (wires: 0 ONE, 1 OUT, 2 X0, 3 K)
    u, v, nx = 3 * r + 4, 3 * r + 5, 3 * r + 6
    nfill = nc - 1 - 3 * rounds
    f = np.arange(nfill, dtype=np.int64)
    last_state = int(nx[-1]) if rounds else 2
    fx_in = np.where(f == 0, last_state, 3 * rounds + 3 + f)
    fx_out = 3 * rounds + 4 + f
    final_state = int(fx_out[-1]) if nfill else last_state
    t_w = np.stack([xw, np.full(rounds, 3), np.zeros(rounds, dtype=np.int64)], axis=1)        # t = x + k + c_i * ONE
    t_c = np.stack([np.zeros(rounds, dtype=np.int64), np.zeros(rounds, dtype=np.int64), 1 + r], axis=1)

    z1 = np.zeros((rounds, 1), dtype=np.int64)
    col = lambda a: a.reshape(-1, 1)

    def matrix(per_round, fill_w, last_w):
        """per_round: the three rows of a round as (wires, coeff ids) pairs of (rounds, k) arrays"""
        lens = [p[0].shape[1] for p in per_round]
        w = np.concatenate([p[0] for p in per_round], axis=1).reshape(-1)
        c = np.concatenate([p[1] for p in per_round], axis=1).reshape(-1)
        row_len = np.concatenate([np.tile(lens, rounds), np.ones(nfill, dtype=np.int64), [len(last_w)]])
        rp = np.zeros(nc + 1, dtype=np.uint32)
        rp[1:] = np.cumsum(row_len)
        w = np.concatenate([w, fill_w, last_w])
        c = np.concatenate([c, np.zeros(nfill + len(last_w), dtype=np.int64)])
        return rp, w.astype(np.uint32), c.astype(np.uint32)

    L = matrix([(t_w, t_c), (col(u), z1), (col(v), z1)], fx_in, [final_state, 3])
    Rm = matrix([(t_w, t_c), (col(u), z1), (t_w, t_c)], fx_in, [0])
    O = matrix([(col(u), z1), (col(v), z1), (col(nx), z1)], fx_out, [1])
    return ConstraintSystem(nb_constraints=nc, nb_wires=nc + 3, coefficients=coeffs, L=L, R=Rm, O=O)
