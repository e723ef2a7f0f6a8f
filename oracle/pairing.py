"""Ate pairings of the two curves on Python ints (oracle; test infrastructure only).

Used only by the restated verifier (backend/groth16/bn254/verify.go:49-148:
curve.MillerLoop :77,:131, curve.FinalExponentiation :141, vk.Precompute
setup.go:353-362; backend/groth16/bls12-377/verify.go is the same file over the
other curve package).  Any correct non-degenerate bilinear map accepts exactly the
same proofs, so this follows the textbook construction: Fp12 = Fp[w]/(modulus of
degree 12), G2 untwisted into E(Fp12), Miller loop with generic line functions,
final exponent (p^12-1)/r.
  BN254:     u^2 = -1, xi = 9 + u = w^6  =>  w^12 = 18 w^6 - 82; loop over 6x+2 with the
             two Frobenius correction steps (optimal ate).
  BLS12-377: u^2 = -5, xi = u = w^6 (gnark-crypto's tower: v^3 = u, w^2 = v)  =>  w^12 = -5;
             twist y^2 = x^3 + 1/u (D type), loop over the seed x = 0x8508c00000000001 (ate
             pairing with T = t - 1 = x), no correction steps.
Like the rest of the oracle the module is curve-generic through the alias package
(oracle.pairing = BN254, oracle.bls12_377.pairing = BLS12-377).
"""
from .curve import NAME, P, R, inv

if NAME == "bn254":
    _MOD = (82, 0, 0, 0, 0, 0, -18, 0, 0, 0, 0, 0)   # w^12 = 18 w^6 - 82
    ATE_LOOP = 29793968203157093288                  # 6x+2, x = 4965661367192848881
    LOG_ATE = 63
    _XI_SHIFT = 9                                    # u = w^6 - 9
else:
    _MOD = (5, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0)      # w^12 = u^2 = -5
    ATE_LOOP = 0x8508c00000000001                    # the curve seed x
    LOG_ATE = 62                                     # bits below the leading one
    _XI_SHIFT = 0                                    # u = w^6


def f12(c0=0, *rest):
    v = [c0 % P] + [c % P for c in rest]
    return tuple(v + [0] * (12 - len(v)))


F12_ONE = f12(1)
F12_ZERO = f12(0)


def f12_add(a, b):
    return tuple((x + y) % P for x, y in zip(a, b))


def f12_sub(a, b):
    return tuple((x - y) % P for x, y in zip(a, b))


def f12_mul(a, b):
    t = [0] * 23
    for i, x in enumerate(a):
        if x:
            for j, y in enumerate(b):
                t[i + j] += x * y
    for k in range(22, 11, -1):          # reduce w^k = w^(k-12) * (-_MOD[6] w^6 - _MOD[0])
        c = t[k]
        if c:
            if _MOD[6]:
                t[k - 6] -= _MOD[6] * c
            t[k - 12] -= _MOD[0] * c
    return tuple(x % P for x in t[:12])


def f12_scalar(a, k):
    return tuple(x * k % P for x in a)


def _deg(p):
    d = len(p) - 1
    while d and p[d] == 0:
        d -= 1
    return d


def f12_inv(a):
    """extended Euclid over Fp[w] modulo the degree-12 modulus."""
    lm, hm = [1] + [0] * 12, [0] * 13
    low = list(a) + [0]
    high = [c % P for c in _MOD] + [1]
    while _deg(low):
        # r = high / low
        dl = _deg(low)
        r = [0] * 13
        tmp = list(high)
        il = inv(low[dl], P)
        for i in range(_deg(tmp) - dl, -1, -1):
            q = tmp[dl + i] * il % P
            r[i] = q
            for c in range(dl + 1):
                tmp[c + i] = (tmp[c + i] - q * low[c]) % P
        nm, new = list(hm), list(high)
        for i in range(13):
            for j in range(13 - i):
                nm[i + j] = (nm[i + j] - lm[i] * r[j]) % P
                new[i + j] = (new[i + j] - low[i] * r[j]) % P
        lm, low, hm, high = nm, new, lm, low
    il = inv(low[0], P)
    return tuple(c * il % P for c in lm[:12])


def f12_pow(a, e):
    out = F12_ONE
    while e:
        if e & 1:
            out = f12_mul(out, a)
        a = f12_mul(a, a)
        e >>= 1
    return out


_W = f12(0, 1)
_W2 = f12_mul(_W, _W)
_W3 = f12_mul(_W2, _W)


def _twist(q):
    """E'(Fp2) -> E(Fp12): (x, y) -> (x w^2, y w^3) with u = w^6 - _XI_SHIFT embedded.  With xi = w^6:
    (y w^3)^2 = y^2 xi and (x w^2)^3 = x^3 xi, so y^2 = x^3 + b/xi (both twists are of D type)."""
    (x0, x1), (y0, y1) = q
    nx = f12(x0 - _XI_SHIFT * x1, 0, 0, 0, 0, 0, x1)
    ny = f12(y0 - _XI_SHIFT * y1, 0, 0, 0, 0, 0, y1)
    return (f12_mul(nx, _W2), f12_mul(ny, _W3))


def _cast(p):
    return (f12(p[0]), f12(p[1]))


def _double(pt):
    x, y = pt
    m = f12_mul(f12_scalar(f12_mul(x, x), 3), f12_inv(f12_scalar(y, 2)))
    nx = f12_sub(f12_mul(m, m), f12_scalar(x, 2))
    ny = f12_sub(f12_mul(m, f12_sub(x, nx)), y)
    return (nx, ny)


def _add(p1, p2):
    if p1 is None:
        return p2
    if p2 is None:
        return p1
    x1, y1 = p1
    x2, y2 = p2
    if x1 == x2:
        return _double(p1) if y1 == y2 else None
    m = f12_mul(f12_sub(y2, y1), f12_inv(f12_sub(x2, x1)))
    nx = f12_sub(f12_sub(f12_mul(m, m), x1), x2)
    ny = f12_sub(f12_mul(m, f12_sub(x1, nx)), y1)
    return (nx, ny)


def _line(p1, p2, t):
    x1, y1 = p1
    x2, y2 = p2
    xt, yt = t
    if x1 != x2:
        m = f12_mul(f12_sub(y2, y1), f12_inv(f12_sub(x2, x1)))
    elif y1 == y2:
        m = f12_mul(f12_scalar(f12_mul(x1, x1), 3), f12_inv(f12_scalar(y1, 2)))
    else:
        return f12_sub(xt, x1)
    return f12_sub(f12_mul(m, f12_sub(xt, x1)), f12_sub(yt, y1))


def miller_loop(p_g1, q_g2):
    """f_{T,Q}(P), T = ATE_LOOP (BN254: with the Frobenius corrections); no final exponentiation."""
    if p_g1 is None or q_g2 is None:
        return F12_ONE
    Q = _twist(q_g2)
    Pt = _cast(p_g1)
    Rp = Q
    f = F12_ONE
    for i in range(LOG_ATE, -1, -1):
        f = f12_mul(f12_mul(f, f), _line(Rp, Rp, Pt))
        Rp = _double(Rp)
        if ATE_LOOP & (1 << i):
            f = f12_mul(f, _line(Rp, Q, Pt))
            Rp = _add(Rp, Q)
    if NAME != "bn254":
        return f
    Q1 = (f12_pow(Q[0], P), f12_pow(Q[1], P))
    nQ2 = (f12_pow(Q1[0], P), f12_sub(F12_ZERO, f12_pow(Q1[1], P)))
    f = f12_mul(f, _line(Rp, Q1, Pt))
    Rp = _add(Rp, Q1)
    f = f12_mul(f, _line(Rp, nQ2, Pt))
    return f


def final_exp(f):
    return f12_pow(f, (P ** 12 - 1) // R)


def pairing_product(pairs):
    """prod e(P_i, Q_i) for [(P_i in G1, Q_i in G2)]."""
    f = F12_ONE
    for p1, q2 in pairs:
        f = f12_mul(f, miller_loop(p1, q2))
    return final_exp(f)
