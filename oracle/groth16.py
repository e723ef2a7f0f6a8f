"""Groth16 BN254 setup / prove / verify restated on Python ints
(oracle; test infrastructure only).

Follows backend/groth16/bn254/{setup.go:85-349,364-446, prove.go:62-399,
verify.go:49-148} for circuits without BSB22 commitments (the BASELINE configs
have none; commitment handling is SURVEY.md 8(f) N3).  r, s and the toxic waste
are injected so proofs are reproducible byte for byte.
"""
import heapq
from . import curve as C
from .curve import R, inv
from .ntt import Domain, bit_reverse_list, compute_h
from .msm import batch_scalar_mul_gen, msm_pippenger
if C.NAME == "bn254":
    from .pairing import pairing_product
else:
    pairing_product = None   # BLS12-377: no restated pairing; proofs are checked on the exponent side (trapdoor)


def filter_heap(slice_, first_index, to_remove):
    """prove.go:331-354 (KATs: backend/groth16/bn254/utils_test.go:17-38)."""
    if not to_remove:
        return slice_
    heap = list(to_remove)
    heapq.heapify(heap)
    out = []
    for i, v in enumerate(slice_):
        if heap and i + first_index == heap[0]:
            while heap and i + first_index == heap[0]:
                heapq.heappop(heap)
            continue
        out.append(v)
    return out


class ToxicWaste:
    def __init__(self, rng):
        def nz():
            while True:
                v = rng.field(R)
                if v:
                    return v
        self.t, self.alpha, self.beta, self.gamma, self.delta = nz(), nz(), nz(), nz(), nz()
        self.gamma_inv = inv(self.gamma, R)
        self.delta_inv = inv(self.delta, R)


def setup_abc(cs, d, tw):
    """setup.go:364-446: A_i(t), B_i(t), C_i(t) via Lagrange basis at t."""
    nw = cs.nb_wires
    A, B, Cc = [0] * nw, [0] * nw, [0] * nw
    t = tw.t
    L = (pow(t, d.n, R) - 1) * inv((t - 1) % R, R) % R * d.n_inv % R
    wi = 1
    for (Le, Re, Oe) in cs.constraints:
        for cf, i in Le:
            A[i] = (A[i] + cf * L) % R
        for cf, i in Re:
            B[i] = (B[i] + cf * L) % R
        for cf, i in Oe:
            Cc[i] = (Cc[i] + cf * L) % R
        # L_{i+1} = w * L_i * (t - w^i) / (t - w^(i+1))
        nwi = wi * d.gen % R
        L = L * d.gen % R * ((t - wi) % R) % R * inv((t - nwi) % R, R) % R
        wi = nwi
    return A, B, Cc


class ProvingKey:
    pass


class VerifyingKey:
    pass


def setup(cs, rng, tw=None):
    """setup.go:85-349 without commitments.  Returns (pk, vk, trapdoor scalars)."""
    tw = tw or ToxicWaste(rng)
    d = Domain(cs.nb_constraints)
    A, B, Cc = setup_abc(cs, d, tw)
    nb_pub = cs.nb_public
    K = [(A[i] * tw.beta + B[i] * tw.alpha + Cc[i]) % R for i in range(cs.nb_wires)]
    vkK = [K[i] * tw.gamma_inv % R for i in range(nb_pub)]
    pkK = [K[i] * tw.delta_inv % R for i in range(nb_pub, cs.nb_wires)]
    zdt = (pow(tw.t, d.n, R) - 1) * tw.delta_inv % R
    Z = []
    for _ in range(d.n):
        Z.append(zdt)
        zdt = zdt * tw.t % R
    pk, vk = ProvingKey(), VerifyingKey()
    pk.infinity_a = [a == 0 for a in A]
    pk.infinity_b = [b == 0 for b in B]
    A_nz = [a for a in A if a]
    B_nz = [b for b in B if b]
    pk.nb_infinity_a = cs.nb_wires - len(A_nz)
    pk.nb_infinity_b = cs.nb_wires - len(B_nz)
    Zbr = bit_reverse_list(Z)[:d.n - 1]          # setup.go:265-267
    g1s = [tw.alpha, tw.beta, tw.delta] + A_nz + B_nz + Zbr + vkK + pkK
    pts = batch_scalar_mul_gen(g1s)
    o = 3
    pk.g1_alpha, pk.g1_beta, pk.g1_delta = pts[0], pts[1], pts[2]
    pk.g1_a = pts[o:o + len(A_nz)]; o += len(A_nz)
    pk.g1_b = pts[o:o + len(B_nz)]; o += len(B_nz)
    pk.g1_z = pts[o:o + len(Zbr)]; o += len(Zbr)
    vk.g1_k = pts[o:o + len(vkK)]; o += len(vkK)
    pk.g1_k = pts[o:o + len(pkK)]; o += len(pkK)
    assert o == len(pts)
    g2 = batch_scalar_mul_gen(B_nz + [tw.beta, tw.delta, tw.gamma], g2=True)
    pk.g2_b = g2[:len(B_nz)]
    pk.g2_beta, pk.g2_delta = g2[len(B_nz)], g2[len(B_nz) + 1]
    vk.g2_delta, vk.g2_gamma = g2[len(B_nz) + 1], g2[len(B_nz) + 2]
    vk.g1_alpha, vk.g2_beta = pk.g1_alpha, pk.g2_beta
    vk.e = pairing_product([(vk.g1_alpha, vk.g2_beta)]) if pairing_product else None   # vk.Precompute setup.go:353-362
    pk.domain = d
    pk.nb_public = nb_pub
    scal = dict(A=A_nz, B=B_nz, Z=Zbr, K=pkK, tw=tw)
    return pk, vk, scal


def prove_parts(cs, pk, w, msm=None):
    """The five raw MSM results + h (what crosses the device boundary).
    prove.go:141-178 (filtering), :202-304 (MSM operands), :356-399 (computeH)."""
    msm = msm or (lambda pts, sc, g2=False: msm_pippenger(pts, sc, c=6, g2=g2))
    a, b, c = cs.abc(w)
    h = compute_h(a, b, c, pk.domain)
    wa = [w[i] for i in range(len(w)) if not pk.infinity_a[i]]
    wb = [w[i] for i in range(len(w)) if not pk.infinity_b[i]]
    wk = filter_heap(w[pk.nb_public:], pk.nb_public, [])
    size_h = pk.domain.n - 1
    return dict(
        h=h, wa=wa, wb=wb, wk=wk,
        ar=msm(pk.g1_a, wa), bs1=msm(pk.g1_b, wb),
        krs_k=msm(pk.g1_k, wk), krs_hz=msm(pk.g1_z, h[:size_h]),
        bs2=msm(pk.g2_b, wb, g2=True),
    )


def assemble(pk, parts, r, s):
    """host epilogue prove.go:180-195, 209-224, 251-279, 297-302."""
    kr = (-r * s) % R
    d_r, d_s, d_kr = (C.g1_mul(pk.g1_delta, k) for k in (r, s, kr))
    ar = C.g1_add(C.g1_add(parts["ar"], pk.g1_alpha), d_r)
    bs1 = C.g1_add(C.g1_add(parts["bs1"], pk.g1_beta), d_s)
    krs = C.g1_add(parts["krs_k"], d_kr)
    krs = C.g1_add(krs, parts["krs_hz"])
    krs = C.g1_add(krs, C.g1_mul(ar, s))
    krs = C.g1_add(krs, C.g1_mul(bs1, r))
    bs = C.g2_add(C.g2_add(parts["bs2"], C.g2_mul(pk.g2_delta, s)), pk.g2_beta)
    return dict(ar=ar, bs=bs, krs=krs)


def prove(cs, pk, w, r, s, msm=None):
    return assemble(pk, prove_parts(cs, pk, w, msm), r, s)


def proof_bytes(proof):
    """raw memory image: Ar (64) || Bs (128) || Krs (64), gnark Montgomery layout."""
    return C.g1_to_bytes(proof["ar"]) + C.g2_to_bytes(proof["bs"]) + C.g1_to_bytes(proof["krs"])


def verify(proof, vk, public_witness):
    """verify.go:49-148 (no commitments): e(Krs,-delta) e(Ar,Bs) e(kSum,-gamma) == e(alpha,beta)."""
    if pairing_product is None:
        raise NotImplementedError("the restated pairing check exists for BN254 only")
    if len(public_witness) != len(vk.g1_k) - 1:
        raise ValueError("invalid witness size")
    for pt, on in ((proof["ar"], C.g1_on_curve), (proof["krs"], C.g1_on_curve), (proof["bs"], C.g2_on_curve)):
        if not on(pt):
            return False
    if C.g2_mul(proof["bs"], R - 1) != C.g2_neg(proof["bs"]):   # subgroup check (prove.go:52-54)
        return False
    ksum = vk.g1_k[0]
    for x, k in zip(public_witness, vk.g1_k[1:]):
        ksum = C.g1_add(ksum, C.g1_mul(k, x))
    lhs = pairing_product([
        (proof["krs"], C.g2_neg(vk.g2_delta)),
        (proof["ar"], proof["bs"]),
        (ksum, C.g2_neg(vk.g2_gamma)),
    ])
    return lhs == vk.e
