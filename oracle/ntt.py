"""Fr FFT domain + computeH on Python ints (oracle; test infrastructure only).

Restates gnark-crypto's fft.Domain as the reference uses it
(backend/groth16/bn254/prove.go:356-399, setup.go:111):
  FFTInverse(x, DIF)          natural in  -> bit-reversed out, scaled by 1/n
  FFT(x, DIT, OnCoset())      bit-reversed in -> natural out; coefficient i
                              (stored at bitrev(i)) is first scaled by g^i
  FFTInverse(x, DIF, OnCoset) natural in -> bit-reversed out, then coefficient
                              i (at bitrev(i)) scaled by g^-i / n
Results are exact field elements, so any correct algorithm is byte-identical
(SURVEY.md F6).
"""
from .curve import R, FR_GEN, FR_ROOT, FR_TWO_ADICITY, inv


def bitrev(i, logn):
    r = 0
    for _ in range(logn):
        r = (r << 1) | (i & 1)
        i >>= 1
    return r


def bit_reverse_list(a):
    """setup.go:690-700 bitReverse (in place permutation, returned as new list)."""
    n = len(a)
    logn = n.bit_length() - 1
    return [a[bitrev(i, logn)] for i in range(n)]


class Domain:
    """fft.NewDomain(m): cardinality = next pow2 >= m (setup.go:111)."""

    def __init__(self, m):
        n = 1
        while n < m:
            n <<= 1
        self.n = n
        self.logn = n.bit_length() - 1
        assert self.logn <= FR_TWO_ADICITY
        self.gen = pow(FR_ROOT, 1 << (FR_TWO_ADICITY - self.logn), R)
        self.gen_inv = inv(self.gen, R)
        self.n_inv = inv(n, R)
        self.coset = FR_GEN
        self.coset_inv = inv(FR_GEN, R)


def _dif(a, w):
    """in-place radix-2 DIF, natural in -> bit-reversed out."""
    n = len(a)
    half = n >> 1
    wstep = w
    while half >= 1:
        for start in range(0, n, 2 * half):
            t = 1
            for j in range(start, start + half):
                u, v = a[j], a[j + half]
                a[j] = (u + v) % R
                a[j + half] = (u - v) * t % R
                t = t * wstep % R
        wstep = wstep * wstep % R
        half >>= 1


def _dit(a, w):
    """in-place radix-2 DIT, bit-reversed in -> natural out."""
    n = len(a)
    logn = n.bit_length() - 1
    ws = [w]
    for _ in range(logn - 1):
        ws.append(ws[-1] * ws[-1] % R)
    half = 1
    lvl = logn - 1
    while half < n:
        wstep = ws[lvl] if logn else 1
        for start in range(0, n, 2 * half):
            t = 1
            for j in range(start, start + half):
                u, v = a[j], a[j + half] * t % R
                a[j] = (u + v) % R
                a[j + half] = (u - v) % R
                t = t * wstep % R
        half <<= 1
        lvl -= 1


def fft_inverse_dif(a, d, coset=False):
    a = list(a)
    assert len(a) == d.n
    if d.n > 1:
        _dif(a, d.gen_inv)
    if coset:
        out = [0] * d.n
        for i in range(d.n):   # coefficient i sits at bitrev(i)
            out[bitrev(i, d.logn)] = a[bitrev(i, d.logn)] * pow(d.coset_inv, i, R) % R * d.n_inv % R
        return out
    return [x * d.n_inv % R for x in a]


def fft_dit(a, d, coset=False):
    a = list(a)
    assert len(a) == d.n
    if coset:
        for i in range(d.n):
            j = bitrev(i, d.logn)
            a[j] = a[j] * pow(d.coset, i, R) % R
    if d.n > 1:
        _dit(a, d.gen)
    return a


def compute_h(a, b, c, d):
    """prove.go:356-399.  Returns h in bit-reversed coefficient order (len n)."""
    pad = d.n - len(a)
    a = list(a) + [0] * pad
    b = list(b) + [0] * pad
    c = list(c) + [0] * pad
    a = fft_inverse_dif(a, d)
    b = fft_inverse_dif(b, d)
    c = fft_inverse_dif(c, d)
    a = fft_dit(a, d, coset=True)
    b = fft_dit(b, d, coset=True)
    c = fft_dit(c, d, coset=True)
    den = inv((pow(d.coset, d.n, R) - 1) % R, R)
    a = [(x * y - z) * den % R for x, y, z in zip(a, b, c)]
    return fft_inverse_dif(a, d, coset=True)


def poly_eval(coeffs, x):
    acc = 0
    for cf in reversed(coeffs):
        acc = (acc * x + cf) % R
    return acc
