"""Multi-scalar multiplication on Python ints (oracle; test infrastructure only).

Restates what the reference obtains from gnark-crypto's G1Jac.MultiExp /
G2Jac.MultiExp (call sites backend/groth16/bn254/prove.go:204,217,237,247,293):
the group element sum_i s_i * P_i, compared in affine form.
Two independent algorithms (double-and-add, bucket method) cross-check each
other in tests/test_oracle_msm.py.
"""
from . import curve as C


def msm_naive(points, scalars, g2=False):
    mul = C.g2_mul if g2 else C.g1_mul
    add = C.g2_add if g2 else C.g1_add
    acc = None
    for p, s in zip(points, scalars):
        acc = add(acc, mul(p, s))
    return acc


def msm_pippenger(points, scalars, c=8, g2=False):
    """unsigned-window bucket method; affine in/out."""
    F = C._Fp2 if g2 else C._Fp
    jadd = lambda a, b: C._jac_add(F, a, b)
    jdbl = lambda a: C._jac_double(F, a)
    ident = (F.one, F.one, F.zero)
    nwin = (254 + c - 1) // c
    total = ident
    for w in range(nwin - 1, -1, -1):
        for _ in range(c):
            total = jdbl(total)
        buckets = [ident] * ((1 << c) - 1)
        for p, s in zip(points, scalars):
            if p is None:
                continue
            d = ((s % C.R) >> (w * c)) & ((1 << c) - 1)
            if d:
                buckets[d - 1] = jadd(buckets[d - 1], C._to_jac(F, p))
        run = ident
        acc = ident
        for bk in reversed(buckets):
            run = jadd(run, bk)
            acc = jadd(acc, run)
        total = jadd(total, acc)
    return C._to_aff(F, total)


def fixed_base_table(gen, g2=False, window=8):
    """[d * 2^(window*j)] gen for d in 1..2^window-1 (jacobian), for fast [k]G."""
    F = C._Fp2 if g2 else C._Fp
    tables = []
    base = C._to_jac(F, gen)
    nwin = (254 + window - 1) // window
    for _ in range(nwin):
        row = [base]
        for _ in range((1 << window) - 2):
            row.append(C._jac_add(F, row[-1], base))
        tables.append(row)
        nb = row[-1]
        nb = C._jac_add(F, nb, base)   # 2^window * base
        base = nb
    return tables


def fixed_base_mul(tables, k, g2=False, window=8):
    F = C._Fp2 if g2 else C._Fp
    k %= C.R
    acc = (F.one, F.one, F.zero)
    j = 0
    mask = (1 << window) - 1
    while k:
        d = k & mask
        if d:
            acc = C._jac_add(F, acc, tables[j][d - 1])
        k >>= window
        j += 1
    return acc


def batch_to_affine(jacs, g2=False):
    F = C._Fp2 if g2 else C._Fp
    zs = [p[2] for p in jacs]
    pref = []
    acc = F.one
    for z in zs:
        pref.append(acc)
        if z != F.zero:
            acc = F.mul(acc, z)
    ai = F.inv(acc)
    out = [None] * len(jacs)
    for i in range(len(jacs) - 1, -1, -1):
        z = zs[i]
        if z == F.zero:
            continue
        zi = F.mul(ai, pref[i])
        ai = F.mul(ai, z)
        zi2 = F.mul(zi, zi)
        out[i] = (F.mul(jacs[i][0], zi2), F.mul(jacs[i][1], F.mul(zi2, zi)))
    return out


_TABLES = {}


def batch_scalar_mul_gen(scalars, g2=False):
    """curve.BatchScalarMultiplicationG1/G2(&gen, scalars) (setup.go:251,320) -> affine."""
    key = bool(g2)
    if key not in _TABLES:
        _TABLES[key] = fixed_base_table(C.G2_GEN if g2 else C.G1_GEN, g2=g2)
    t = _TABLES[key]
    return batch_to_affine([fixed_base_mul(t, s, g2=g2) for s in scalars], g2=g2)
