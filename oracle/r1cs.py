"""Minimal R1CS container + the synthetic circuits of the BASELINE configs
(oracle; test infrastructure only).

Wire numbering follows constraint/core.go:227-248: [ONE, public..., secret...,
internal...].  A constraint is (L, R, O), each a list of (coeff, wire) with
L.w * R.w == O.w (constraint/bn254/solver.go:540-628 accumulates exactly these
three inner products into a[c], b[c], c[c]).
"""
from .curve import R


class R1CS:
    def __init__(self, nb_public, nb_secret):
        # nb_public includes the ONE wire (GetNbPublicVariables)
        self.nb_public = nb_public
        self.nb_secret = nb_secret
        self.nb_internal = 0
        self.constraints = []
        self.commitment_info = []      # constraint.Groth16Commitments (oracle.commitment.Groth16Commitment)

    @property
    def nb_wires(self):
        return self.nb_public + self.nb_secret + self.nb_internal

    @property
    def nb_constraints(self):
        return len(self.constraints)

    def new_internal(self):
        idx = self.nb_wires
        self.nb_internal += 1
        return idx

    def add(self, L, Rr, O):
        self.constraints.append((L, Rr, O))

    def abc(self, w):
        """per-constraint L.w, R.w, O.w; raises if unsatisfied."""
        def dot(le):
            return sum(cf * w[i] for cf, i in le) % R
        a, b, c = [], [], []
        for k, (L, Rr, O) in enumerate(self.constraints):
            x, y, z = dot(L), dot(Rr), dot(O)
            if x * y % R != z:
                raise ValueError(f"constraint #{k} is not satisfied")
            a.append(x)
            b.append(y)
            c.append(z)
        return a, b, c


def cubic_circuit():
    """examples/cubic/cubic.go:32-36: x^3 + x + 5 == y.  Wires: 0=ONE 1=Y 2=X 3=x^2 4=x^3."""
    cs = R1CS(nb_public=2, nb_secret=1)
    v1 = cs.new_internal()
    v2 = cs.new_internal()
    cs.add([(1, 2)], [(1, 2)], [(1, v1)])
    cs.add([(1, v1)], [(1, 2)], [(1, v2)])
    cs.add([(1, v2), (1, 2), (5, 0)], [(1, 0)], [(1, 1)])
    return cs


def cubic_witness(x, y):
    x %= R
    return [1, y % R, x, x * x % R, x * x * x % R]


def mimc_chain(nb_constraints, rng, power=5):
    """MiMC-shaped chain (std/hash/mimc/encrypt.go:124-128,147-153, pow5 rounds; power=17 gives the BLS12-377
    round of encrypt.go:137-143: four squarings and one product, 5 constraints per round):
    x <- (x + k + c_i)^5, 3 constraints per round, then one assertion that binds
    the public output; filler squarings bring the count to exactly nb_constraints.
    Wires: 0=ONE 1=OUT(public) 2=X0(secret) 3=K(secret).  Returns (cs, consts)."""
    cs = R1CS(nb_public=2, nb_secret=2)
    consts = []
    x = [(1, 2)]            # linear expression for the running state
    assert power in (5, 17)
    per_round = 3 if power == 5 else 5
    rounds = (nb_constraints - 1) // per_round
    for _ in range(rounds):
        ci = rng.field(R)
        consts.append(ci)
        t = x + [(1, 3), (ci, 0)]
        u = cs.new_internal()
        cs.add(t, t, [(1, u)])
        v = cs.new_internal()
        cs.add([(1, u)], [(1, u)], [(1, v)])
        if power == 17:
            for _ in range(2):
                v2 = cs.new_internal()
                cs.add([(1, v)], [(1, v)], [(1, v2)])
                v = v2
        nx = cs.new_internal()
        cs.add([(1, v)], t, [(1, nx)])
        x = [(1, nx)]
    while cs.nb_constraints < nb_constraints - 1:
        nx = cs.new_internal()
        cs.add(x, x, [(1, nx)])
        x = [(1, nx)]
    cs.add(x + [(1, 3)], [(1, 0)], [(1, 1)])
    return cs, consts


def mimc_chain_witness(cs, x0, k):
    """solve the chain: every constraint defines at most one new wire (its O)."""
    w = [0] * cs.nb_wires
    w[0], w[2], w[3] = 1, x0 % R, k % R
    def dot(le):
        return sum(cf * w[i] for cf, i in le) % R
    for L, Rr, O in cs.constraints:
        (cf, o), = O
        w[o] = dot(L) * dot(Rr) % R
    return w


def committed_chain(nb_constraints, rng, nb_commitments=1, nb_private_committed=5):
    """MiMC-shaped chain with BSB22 commitments (the shape api.Commit gives a circuit: constraint/commitment.go:9-14,
    backend/groth16/bn254/commitment_test.go:33-49,129-149).  Wires: 0=ONE 1=OUT 2=P (public) 3=X0 4=K (secret).
    The chain runs in nb_commitments + 1 phases; after phase j a commitment wire cm_j (an internal wire no
    constraint defines -- the bsb22 hint does) is created and added to the round key of the later phases.
    Commitment 0 commits to P (public), X0, K and up to nb_private_committed internal wires of phase 0;
    commitment j > 0 commits to P (public), cm_{j-1} (a commitment wire, exercising the index translation
    of GetPublicAndCommitmentCommitted) and internal wires of phase j.  The last commitment of an odd count > 2
    commits to nothing privately (empty basis).  Returns the R1CS."""
    from .commitment import Groth16Commitment
    cs = R1CS(nb_public=3, nb_secret=2)
    phases = nb_commitments + 1
    per_phase = max(1, (nb_constraints - 1) // (3 * phases))
    x = [(1, 3)]
    key = [(1, 4)]
    prev_cm = None
    for ph in range(phases):
        internals = []
        for _ in range(per_phase):
            if cs.nb_constraints + 3 > nb_constraints - 1:
                break
            ci = rng.field(R)
            t = x + key + [(ci, 0)]
            u = cs.new_internal()
            cs.add(t, t, [(1, u)])
            v = cs.new_internal()
            cs.add([(1, u)], [(1, u)], [(1, v)])
            nx = cs.new_internal()
            cs.add([(1, v)], t, [(1, nx)])
            internals += [u, v, nx]
            x = [(1, nx)]
        if ph < nb_commitments:
            if ph == 0:
                pub, priv = [2], [3, 4] + internals[:nb_private_committed]
                nb_pub_committed = 1
            else:
                pub, priv = [2, prev_cm], internals[1:1 + nb_private_committed]
                nb_pub_committed = 1
                if nb_commitments > 2 and ph == nb_commitments - 1 and nb_commitments % 2 == 1:
                    priv = []
            cm = cs.new_internal()
            cs.commitment_info.append(Groth16Commitment(pub, sorted(priv), cm, nb_pub_committed))
            key = key + [(1, cm)]
            prev_cm = cm
    while cs.nb_constraints < nb_constraints - 1:
        nx = cs.new_internal()
        cs.add(x + key[1:], x, [(1, nx)])
        x = [(1, nx)]
    cs.add(x + key, [(1, 0)], [(1, 1)])
    return cs


def committed_chain_assignment(cs, pk, p, x0, k, solve):
    """input wires [ONE, OUT, P, X0, K]: OUT is what the chain evaluates to, found by solving everything but the
    final assertion once (no commitment hashes OUT, so there is no circularity)."""
    probe = [1, 0, p % R, x0 % R, k % R]
    saved = cs.constraints
    cs.constraints = saved[:-1]
    try:
        w, _, _ = solve(cs, pk, probe)
    finally:
        cs.constraints = saved
    out = sum(cf * w[j] for cf, j in saved[-1][0]) % R
    return [1, out, p % R, x0 % R, k % R]
