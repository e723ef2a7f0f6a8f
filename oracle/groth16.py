"""Groth16 BN254 setup / prove / verify restated on Python ints
(oracle; test infrastructure only).

Follows backend/groth16/bn254/{setup.go:85-349,364-446, prove.go:62-399,
verify.go:49-148}, BSB22 commitments included (the pedersen /
hash-to-field parts live in commitment.py).  r, s, the toxic waste and the
Pedersen secrets are injected so proofs are reproducible byte for byte.
"""
import heapq
from . import curve as C
from .curve import R, inv
from .ntt import Domain, bit_reverse_list, compute_h
from .msm import batch_scalar_mul_gen, msm_pippenger
from . import commitment as CM
from .pairing import pairing_product      # curve-generic like this module (BN254 optimal ate / BLS12-377 ate)


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
    """setup.go:85-349.  Returns (pk, vk, trapdoor scalars).  Wires that a BSB22 commitment commits to privately
    leave pk.G1.K for that commitment's Pedersen basis (scaled by 1/gamma), commitment wires join vk.G1.K
    (setup.go:104-108,145-196,278-307)."""
    tw = tw or ToxicWaste(rng)
    d = Domain(cs.nb_constraints)
    A, B, Cc = setup_abc(cs, d, tw)
    nb_pub = cs.nb_public
    info = getattr(cs, "commitment_info", [])
    commitment_wires = CM.commitment_indexes(info)
    private_committed = [c.private_committed for c in info]
    vkK, pkK, ckK = [], [], [[] for _ in info]
    cI = [0] * len(info)
    nb_commitments_seen = 0
    for i in range(cs.nb_wires):                       # setup.go:164-196
        k = (A[i] * tw.beta + B[i] * tw.alpha + Cc[i]) % R
        commitment, is_commitment, is_public = -1, False, i < nb_pub
        if not is_public:
            if nb_commitments_seen < len(commitment_wires) and commitment_wires[nb_commitments_seen] == i:
                is_commitment = True
                nb_commitments_seen += 1
            for j in range(len(info)):
                if cI[j] < len(private_committed[j]) and private_committed[j][cI[j]] == i:
                    commitment = j
                    break
        if is_public or is_commitment:
            vkK.append(k * tw.gamma_inv % R)
        elif commitment != -1:
            ckK[commitment].append(k * tw.gamma_inv % R)
            cI[commitment] += 1
        else:
            pkK.append(k * tw.delta_inv % R)
    assert len(vkK) == nb_pub + len(info)
    assert len(pkK) == cs.nb_wires - nb_pub - len(info) - sum(len(p) for p in private_committed)
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
    ck_flat = [k for row in ckK for k in row]
    g1s = [tw.alpha, tw.beta, tw.delta] + A_nz + B_nz + Zbr + vkK + pkK + ck_flat
    pts = batch_scalar_mul_gen(g1s)
    o = 3
    pk.g1_alpha, pk.g1_beta, pk.g1_delta = pts[0], pts[1], pts[2]
    pk.g1_a = pts[o:o + len(A_nz)]; o += len(A_nz)
    pk.g1_b = pts[o:o + len(B_nz)]; o += len(B_nz)
    pk.g1_z = pts[o:o + len(Zbr)]; o += len(Zbr)
    vk.g1_k = pts[o:o + len(vkK)]; o += len(vkK)
    pk.g1_k = pts[o:o + len(pkK)]; o += len(pkK)
    bases = []
    for row in ckK:
        bases.append(pts[o:o + len(row)]); o += len(row)
    assert o == len(pts)
    # Commitment setup (setup.go:278-307): one G2 point shared by all keys, one sigma per key
    pk.commitment_keys, vk.commitment_keys, sigmas = [], [], []
    if info:
        c_g2 = C.g2_mul(C.G2_GEN, rng.field(R) or 1)          # curve.RandomOnG2
        for i, basis in enumerate(bases):
            sigma = rng.field(R) or 1
            # BasisExpSigma through the trapdoor (basis_j = [ckK_j] G): one batched fixed-base call instead of
            # len(basis) variable-base multiplications; CM.pedersen_setup is the literal form (tests compare both)
            exp_sigma = batch_scalar_mul_gen([k * sigma % R for k in ckK[i]]) if basis else []
            pk.commitment_keys.append(CM.PedersenPK(list(basis), exp_sigma))
            vk.commitment_keys.append(CM.PedersenVK(c_g2, C.g2_mul(c_g2, (-sigma) % R)))
            sigmas.append(sigma)
    vk.public_and_commitment_committed = CM.get_public_and_commitment_committed(info, commitment_wires, nb_pub)
    g2 = batch_scalar_mul_gen(B_nz + [tw.beta, tw.delta, tw.gamma], g2=True)
    pk.g2_b = g2[:len(B_nz)]
    pk.g2_beta, pk.g2_delta = g2[len(B_nz)], g2[len(B_nz) + 1]
    vk.g2_delta, vk.g2_gamma = g2[len(B_nz) + 1], g2[len(B_nz) + 2]
    vk.g1_alpha, vk.g2_beta = pk.g1_alpha, pk.g2_beta
    vk.e = pairing_product([(vk.g1_alpha, vk.g2_beta)])   # vk.Precompute setup.go:353-362
    pk.domain = d
    pk.nb_public = nb_pub
    scal = dict(A=A_nz, B=B_nz, Z=Zbr, K=pkK, tw=tw, CK=ckK, sigma=sigmas)
    return pk, vk, scal


def k_wire_indices(cs):
    """wire indices that feed the K MSM: W[nbPublic:] minus privately committed and commitment wires
    (prove.go:241-246: filterHeap(wireValues[nbPublic:], nbPublic, ConcatAll(toRemove...)))."""
    info = getattr(cs, "commitment_info", [])
    to_remove = [w for c in info for w in c.private_committed] + CM.commitment_indexes(info)
    return filter_heap(list(range(cs.nb_public, cs.nb_wires)), cs.nb_public, to_remove)


def solve(cs, pk, assignment, msm=None, hasher=None, hint_fn=None):
    """r1cs.Solve with the bsb22 hint override of prove.go:81-109.  `assignment` holds the input wires (public and
    secret; ONE included); internal wires are derived constraint by constraint (each constraint of the synthetic
    circuits defines its single O wire), commitment wires by the hint.  Returns (w, commitments, private values).
    `hint_fn(i, hashed, committed) -> int` replaces the hint body (solver.OverrideHint): that is how the tests put
    the device implementation under the same solver."""
    info = getattr(cs, "commitment_info", [])
    by_wire = {c.commitment_index: i for i, c in enumerate(info)}
    nb_in = cs.nb_public + cs.nb_secret
    w = list(assignment[:nb_in]) + [None] * (cs.nb_wires - nb_in)
    commitments, private_values = [None] * len(info), [None] * len(info)

    def hint(i):
        c = info[i]
        hashed = [w[j] for j in c.public_and_commitment_committed]
        private_values[i] = [w[j] for j in c.private_committed]
        assert None not in hashed and None not in private_values[i], "committed wire not solved yet"
        if hint_fn is not None:
            commitments[i] = "external"
            w[c.commitment_index] = hint_fn(i, hashed, private_values[i]) % R
            return
        commitments[i] = CM.pedersen_commit(pk.commitment_keys[i].basis, private_values[i], msm)
        w[c.commitment_index] = CM.solve_commitment_wire(commitments[i], hashed, hasher)

    def dot(le):
        acc = 0
        for cf, j in le:
            if w[j] is None:
                hint(by_wire[j])          # KeyError = a wire no earlier constraint and no hint defines
            acc += cf * w[j]
        return acc % R

    for L, Rr, O in cs.constraints:
        x, y = dot(L), dot(Rr)
        if len(O) == 1 and w[O[0][1]] is None and O[0][1] not in by_wire:
            cf, o = O[0]
            w[o] = x * y % R * inv(cf, R) % R
        elif x * y % R != dot(O):
            raise ValueError("constraint is not satisfied")
    for i in range(len(info)):            # a commitment no constraint reads is still part of the proof
        if commitments[i] is None:
            hint(i)
    return w, commitments, private_values


def prove_commitment_pok(cs, pk, w, private_values, msm=None):
    """prove.go:117-135: per-key proofs of knowledge folded with the G16-BSB22 challenge."""
    info = getattr(cs, "commitment_info", [])
    poks = [CM.pedersen_commit(pk.commitment_keys[i].basis_exp_sigma, private_values[i], msm) for i in range(len(info))]
    challenge = CM.fold_challenge([w[c.commitment_index] for c in info])
    return CM.fold(poks, challenge)


def prove_parts(cs, pk, w, msm=None):
    """The five raw MSM results + h (what crosses the device boundary).
    prove.go:141-178 (filtering), :202-304 (MSM operands), :356-399 (computeH)."""
    msm = msm or (lambda pts, sc, g2=False: msm_pippenger(pts, sc, c=6, g2=g2))
    a, b, c = cs.abc(w)
    h = compute_h(a, b, c, pk.domain)
    wa = [w[i] for i in range(len(w)) if not pk.infinity_a[i]]
    wb = [w[i] for i in range(len(w)) if not pk.infinity_b[i]]
    wk = [w[i] for i in k_wire_indices(cs)]
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


def prove_full(cs, pk, assignment, r, s, msm=None):
    """Prove from the input assignment (prove.go:62-325), commitments and their folded proof of knowledge included."""
    w, commitments, private_values = solve(cs, pk, assignment, msm)
    proof = prove(cs, pk, w, r, s, msm)
    proof["commitments"] = commitments
    proof["commitment_pok"] = prove_commitment_pok(cs, pk, w, private_values, msm) if commitments else None
    return proof, w


def proof_bytes(proof):
    """raw memory image: Ar (64) || Bs (128) || Krs (64), gnark Montgomery layout."""
    return C.g1_to_bytes(proof["ar"]) + C.g2_to_bytes(proof["bs"]) + C.g1_to_bytes(proof["krs"])


def verify(proof, vk, public_witness, hasher=None):
    """verify.go:49-148: e(Krs,-delta) e(Ar,Bs) e(kSum,-gamma) == e(alpha,beta), where kSum also takes the BSB22
    commitments and the commitment wires re-derived by hashing (verify.go:82-126)."""
    pacc = getattr(vk, "public_and_commitment_committed", [])
    nb_public_vars = len(vk.g1_k) - len(pacc)
    if len(public_witness) != nb_public_vars - 1:
        raise ValueError("invalid witness size")
    commitments = proof.get("commitments", [])
    if len(commitments) != len(pacc):
        raise ValueError("invalid proof: wrong number of commitments")
    for pt, on in ((proof["ar"], C.g1_on_curve), (proof["krs"], C.g1_on_curve), (proof["bs"], C.g2_on_curve)):
        if not on(pt):
            return False
    if C.g2_mul(proof["bs"], R - 1) != C.g2_neg(proof["bs"]):   # subgroup check (prove.go:52-54)
        return False
    public_witness = list(public_witness)
    commitment_values = []
    for i, idxs in enumerate(pacc):                               # solveCommitmentWire
        res = CM.solve_commitment_wire(commitments[i], [public_witness[j - 1] for j in idxs], hasher)
        public_witness.append(res)
        commitment_values.append(res)
    if getattr(vk, "commitment_keys", []):
        challenge = CM.fold_challenge(commitment_values)
        if not CM.batch_verify_multi_vk(vk.commitment_keys, commitments, proof.get("commitment_pok"), challenge,
                                        pairing_product):
            return False
    ksum = vk.g1_k[0]
    for x, k in zip(public_witness, vk.g1_k[1:]):
        ksum = C.g1_add(ksum, C.g1_mul(k, x))
    for cm in commitments:
        ksum = C.g1_add(ksum, cm)
    lhs = pairing_product([
        (proof["krs"], C.g2_neg(vk.g2_delta)),
        (proof["ar"], proof["bs"]),
        (ksum, C.g2_neg(vk.g2_gamma)),
    ])
    return lhs == vk.e
