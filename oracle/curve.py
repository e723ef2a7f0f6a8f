"""Fields and groups of the Groth16 curves on Python big ints (oracle; test infrastructure only).

Restates the arithmetic the reference gets from gnark-crypto (not vendored,
go.mod:11): fr/fp elements, E2, G1/G2 affine+Jacobian.  Memory layout follows
gnark-crypto: an element is 4 (Fr, BN254 Fp) or 6 (BLS12-377 Fp) little-endian
uint64 limbs in Montgomery form, R = 2^(64 limbs);
G1Affine = X||Y, G2Affine = X.A0||X.A1||Y.A0||Y.A1; the point at infinity is (0,0).

The module is curve-generic: imported as ``oracle.curve`` (or its alias
``oracle.bn254``) it is BN254; imported through the alias package
``oracle.bls12_377`` (whose __path__ is this directory) every oracle module --
curve, ntt, msm, r1cs, groth16 -- is instantiated a second time for BLS12-377
(BASELINE.json config 5).
"""

NAME = "bls12_377" if (__package__ or "").endswith("bls12_377") else "bn254"

if NAME == "bn254":
    P = 21888242871839275222246405745257275088696311157297823662689037894645226208583
    R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
    FP_BYTES = 32
    NONRESIDUE = 1      # Fp2 = Fp[u]/(u^2 + 1)
    B1 = 3              # G1: y^2 = x^3 + 3
    G1_GEN = (1, 2)
    # G2 generator (EIP-197 / gnark-crypto bn254.Generators)
    G2_GEN = (
        (10857046999023057135944570762232829481370756359578518086990519993285655852781,
         11559732032986387107991004021392285783925812861821192530917403151452391805634),
        (8495653923123431417604973247489272438418190587263600148770280649306958101930,
         4082367875863433681332203403145435568316851327593401208105741076214120093531),
    )
    FR_GEN = 5          # fft.Domain FrMultiplicativeGen (coset shift)
    FR_TWO_ADICITY = 28
else:
    # x = 0x8508c00000000001; r = x^4 - x^2 + 1; p = (x-1)^2 r / 3 + x  (checked in tests/test_oracle_bls12_377.py)
    P = 258664426012969094010652733694893533536393512754914660539884262666720468348340822774968888139573360124440321458177
    R = 8444461749428370424248824938781546531375899335154063827935233455917409239041
    FP_BYTES = 48
    NONRESIDUE = 5      # Fp2 = Fp[u]/(u^2 + 5)  (gnark-crypto bls12-377 E2)
    B1 = 1              # G1: y^2 = x^3 + 1
    # gnark-crypto bls12-377 G1 generator
    G1_GEN = (81937999373150964239938255573465948239988671502647976594219695644855304257327692006745978603320413799295628339695,
              241266749859715473739788878240585681733927191168601896383759122102112907357779751001206799952863815012735208165030)
    # harness generator of the order-r subgroup of the twist y^2 = x^3 + 1/u: cofactor x (the twist point with x = 2).
    # (gnark-crypto's own G2 generator is not needed on the proving path: every G2 operand comes from the key.)
    G2_GEN = (
        (39292833563790338514455678255839969442444299076493345799525535236324569704972737101027043002275594504529645125033,
         97668274349181098911216378040700666521757961257997861327997265570326738925466145318868002777904267769221513117576),
        (245994257517657523171405884474647188067285204768246772529216161539930069107591277111081140518594262108675661622819,
         174231680960632680395570731097190109725774571769655017475028422391967989708646134812133505559105641519841883619409),
    )
    FR_GEN = 22         # fft.Domain FrMultiplicativeGen (coset shift)
    FR_TWO_ADICITY = 47
FR_BYTES = 32
G1_BYTES = 2 * FP_BYTES
G2_BYTES = 4 * FP_BYTES
MONT_R = 1 << 256                 # Fr Montgomery radix
MONT_R_FP = 1 << (8 * FP_BYTES)   # Fp Montgomery radix
# primitive 2^FR_TWO_ADICITY-th root of unity of Fr: FR_GEN^((r-1)/2^adicity) (gnark-crypto's fft.Domain generator;
# for BLS12-377 this is 8065159656716812877374967518403273466521432693661810619979959746626482506078)
FR_ROOT = pow(FR_GEN, (R - 1) >> FR_TWO_ADICITY, R)
FR_ROOT_2_28 = pow(FR_ROOT, 1 << (FR_TWO_ADICITY - 28), R)


def inv(a, m):
    return pow(a, -1, m)


# ---------------------------------------------------------------- Fp2
def f2_add(a, b):
    return ((a[0] + b[0]) % P, (a[1] + b[1]) % P)


def f2_sub(a, b):
    return ((a[0] - b[0]) % P, (a[1] - b[1]) % P)


def f2_neg(a):
    return ((-a[0]) % P, (-a[1]) % P)


def f2_mul(a, b):
    return ((a[0] * b[0] - NONRESIDUE * a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)


def f2_sqr(a):
    return f2_mul(a, a)


def f2_scalar(a, k):
    return (a[0] * k % P, a[1] * k % P)


def f2_inv(a):
    d = inv((a[0] * a[0] + NONRESIDUE * a[1] * a[1]) % P, P)
    return (a[0] * d % P, (-a[1]) * d % P)


F2_ZERO = (0, 0)
F2_ONE = (1, 0)
# twist coefficient: BN254 b' = 3/(9+u) (D-twist); BLS12-377 b' = 1/u
B2 = f2_mul((3, 0), f2_inv((9, 1))) if NAME == "bn254" else f2_inv((0, 1))


class _Fp:
    zero = 0
    one = 1
    add = staticmethod(lambda a, b: (a + b) % P)
    sub = staticmethod(lambda a, b: (a - b) % P)
    mul = staticmethod(lambda a, b: a * b % P)
    neg = staticmethod(lambda a: (-a) % P)
    inv = staticmethod(lambda a: inv(a, P))
    b = B1


class _Fp2:
    zero = F2_ZERO
    one = F2_ONE
    add = staticmethod(f2_add)
    sub = staticmethod(f2_sub)
    mul = staticmethod(f2_mul)
    neg = staticmethod(f2_neg)
    inv = staticmethod(f2_inv)
    b = B2


# ---------------------------------------------------------------- curve ops
# Affine points are (x, y) or None for infinity.  Jacobian points are (X,Y,Z).
def _is_on_curve(F, pt):
    if pt is None:
        return True
    x, y = pt
    return F.mul(y, y) == F.add(F.mul(F.mul(x, x), x), F.b)


def _aff_neg(F, pt):
    return None if pt is None else (pt[0], F.neg(pt[1]))


def _jac_double(F, p):
    X, Y, Z = p
    if Z == F.zero:
        return p
    A = F.mul(X, X)
    Bq = F.mul(Y, Y)
    C = F.mul(Bq, Bq)
    t = F.add(X, Bq)
    D = F.sub(F.sub(F.mul(t, t), A), C)
    D = F.add(D, D)
    E = F.add(F.add(A, A), A)
    Fq = F.mul(E, E)
    X3 = F.sub(Fq, F.add(D, D))
    C8 = F.add(C, C)
    C8 = F.add(C8, C8)
    C8 = F.add(C8, C8)
    Y3 = F.sub(F.mul(E, F.sub(D, X3)), C8)
    Z3 = F.mul(Y, Z)
    Z3 = F.add(Z3, Z3)
    return (X3, Y3, Z3)


def _jac_add(F, p, q):
    if p[2] == F.zero:
        return q
    if q[2] == F.zero:
        return p
    X1, Y1, Z1 = p
    X2, Y2, Z2 = q
    Z1Z1 = F.mul(Z1, Z1)
    Z2Z2 = F.mul(Z2, Z2)
    U1 = F.mul(X1, Z2Z2)
    U2 = F.mul(X2, Z1Z1)
    S1 = F.mul(F.mul(Y1, Z2), Z2Z2)
    S2 = F.mul(F.mul(Y2, Z1), Z1Z1)
    if U1 == U2:
        if S1 == S2:
            return _jac_double(F, p)
        return (F.one, F.one, F.zero)
    H = F.sub(U2, U1)
    Rr = F.sub(S2, S1)
    HH = F.mul(H, H)
    HHH = F.mul(H, HH)
    V = F.mul(U1, HH)
    X3 = F.sub(F.sub(F.mul(Rr, Rr), HHH), F.add(V, V))
    Y3 = F.sub(F.mul(Rr, F.sub(V, X3)), F.mul(S1, HHH))
    Z3 = F.mul(F.mul(Z1, Z2), H)
    return (X3, Y3, Z3)


def _to_jac(F, pt):
    return (F.one, F.one, F.zero) if pt is None else (pt[0], pt[1], F.one)


def _to_aff(F, p):
    if p[2] == F.zero:
        return None
    zi = F.inv(p[2])
    zi2 = F.mul(zi, zi)
    return (F.mul(p[0], zi2), F.mul(p[1], F.mul(zi2, zi)))


def _scalar_mul(F, pt, k):
    """double-and-add, affine in / affine out; k reduced mod r."""
    k %= R
    acc = (F.one, F.one, F.zero)
    base = _to_jac(F, pt)
    while k:
        if k & 1:
            acc = _jac_add(F, acc, base)
        base = _jac_double(F, base)
        k >>= 1
    return _to_aff(F, acc)


def _aff_add(F, a, b):
    return _to_aff(F, _jac_add(F, _to_jac(F, a), _to_jac(F, b)))


# G1
def g1_on_curve(pt): return _is_on_curve(_Fp, pt)
def g1_neg(pt): return _aff_neg(_Fp, pt)
def g1_add(a, b): return _aff_add(_Fp, a, b)
def g1_mul(pt, k): return _scalar_mul(_Fp, pt, k)
def g1_jac_add(p, q): return _jac_add(_Fp, p, q)
def g1_jac_double(p): return _jac_double(_Fp, p)
def g1_to_jac(pt): return _to_jac(_Fp, pt)
def g1_to_aff(p): return _to_aff(_Fp, p)


# G2
def g2_on_curve(pt): return _is_on_curve(_Fp2, pt)
def g2_neg(pt): return _aff_neg(_Fp2, pt)
def g2_add(a, b): return _aff_add(_Fp2, a, b)
def g2_mul(pt, k): return _scalar_mul(_Fp2, pt, k)
def g2_jac_add(p, q): return _jac_add(_Fp2, p, q)
def g2_jac_double(p): return _jac_double(_Fp2, p)
def g2_to_jac(pt): return _to_jac(_Fp2, pt)
def g2_to_aff(p): return _to_aff(_Fp2, p)


# ---------------------------------------------------------------- gnark memory layout
_RINV_R = inv(MONT_R, R)
_RINV_P = inv(MONT_R_FP, P)


def fr_to_bytes(v): return ((v % R) * MONT_R % R).to_bytes(32, "little")
def fr_from_bytes(b): return int.from_bytes(b, "little") * _RINV_R % R
def fp_to_bytes(v): return ((v % P) * MONT_R_FP % P).to_bytes(FP_BYTES, "little")
def fp_from_bytes(b): return int.from_bytes(b, "little") * _RINV_P % P


def fr_vec_to_bytes(vs):
    return b"".join(fr_to_bytes(v) for v in vs)


def fr_vec_from_bytes(buf):
    return [fr_from_bytes(buf[i:i + 32]) for i in range(0, len(buf), 32)]


def g1_to_bytes(pt):
    if pt is None:
        return bytes(G1_BYTES)
    return fp_to_bytes(pt[0]) + fp_to_bytes(pt[1])


def g1_from_bytes(b):
    if b == bytes(G1_BYTES):
        return None
    return (fp_from_bytes(b[:FP_BYTES]), fp_from_bytes(b[FP_BYTES:G1_BYTES]))


def g2_to_bytes(pt):
    if pt is None:
        return bytes(G2_BYTES)
    (x0, x1), (y0, y1) = pt
    return fp_to_bytes(x0) + fp_to_bytes(x1) + fp_to_bytes(y0) + fp_to_bytes(y1)


def g2_from_bytes(b):
    if b == bytes(G2_BYTES):
        return None
    v = [fp_from_bytes(b[i:i + FP_BYTES]) for i in range(0, G2_BYTES, FP_BYTES)]
    return ((v[0], v[1]), (v[2], v[3]))


def g1_vec_to_bytes(pts): return b"".join(g1_to_bytes(p) for p in pts)
def g2_vec_to_bytes(pts): return b"".join(g2_to_bytes(p) for p in pts)
