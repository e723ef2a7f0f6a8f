"""GPU parity at the sizes BASELINE.json's metric is quoted on ("the metric sizes have no byte-level evidence"):
G1 MSM at 2^24 and 2^26 (caller-supplied and resident bases), single transforms at 2^20 / 2^24 / 2^26, computeH at
2^20 / 2^24 / 2^26, and the BN254 prove at 2^22 constraints.  Inputs are generated on the device (nothing this large is
shipped), expectations come from the oracle: exponent-side identities, oracle/c's multithreaded FFT, and -- where even
that is too slow -- closed forms (sparse polynomials whose quotient is known analytically)."""
import numpy as np
import pytest

from gnark_icicle_b200 import lib
from oracle import bn254 as C
from oracle import cpu as ocpu

from helpers import ChainCase, assert_limbs_equal, bitrev_perm, g1_gen_arr

pytestmark = pytest.mark.gpu
R = C.R
FR_TOP_MASK = 0x0FFFFFFFFFFFFFFF


def _rand_fr(torch, n, gen):
    x = torch.randint(-(2 ** 63), 2 ** 63 - 1, (n, 4), dtype=torch.int64, device="cuda:0", generator=gen)
    x[:, 3] &= FR_TOP_MASK          # any 252-bit value is a valid Montgomery image of some Fr element
    return x


def _host(t):
    return t.cpu().numpy().view(np.uint64)


@pytest.mark.parametrize("logn", [24, 26])
def test_msm_g1_metric_sizes_by_trapdoor(ctx, logn):
    """config 2 at the metric size and at the top of the sweep: P_i = [k_i]G from the fixed-base kernel, result must be
    [sum s_i k_i]G (oracle/c dot product + double-and-add) for the per-window path and for resident bases with tables."""
    import torch
    n = 1 << logn
    gen = torch.Generator(device="cuda:0")
    gen.manual_seed(2400 + logn)
    k, s = _rand_fr(torch, n, gen), _rand_fr(torch, n, gen)
    bases = torch.empty((n, 8), dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    ctx.g1_fixed_base_mul_dev(k.data_ptr(), n, bases.data_ptr())
    want = ocpu.g1_mul(g1_gen_arr(), ocpu.fr_dot(_host(s), _host(k)))
    got = ctx.msm_g1_dev(s.data_ptr(), bases.data_ptr(), n)
    assert_limbs_equal(got, want, f"g16_msm_g1_dev 2^{logn} vs trapdoor")
    h = ctx.bases_upload_dev(bases.data_ptr(), n)
    del bases
    try:
        got = ctx.msm_resident_dev(h, s.data_ptr())
        assert_limbs_equal(got, want, f"g16_msm_resident_dev 2^{logn} vs trapdoor")
    finally:
        ctx.bases_free(h)


@pytest.mark.parametrize("logn", [20, 24])
def test_transforms_match_oracle_fft_at_large_sizes(ctx, logn):
    """the three transform shapes computeH is made of (prove.go:372-378,396), full compare with oracle/c's FFT"""
    import torch
    n = 1 << logn
    gen = torch.Generator(device="cuda:0")
    gen.manual_seed(3000 + logn)
    x = _rand_fr(torch, n, gen)
    xh = _host(x)
    for name, kw, mode in (("FFTInverse(DIF)", dict(inverse=True, coset=False, decimation=lib.DIF), ocpu.FFT_INV_DIF),
                           ("FFT(DIT, OnCoset)", dict(inverse=False, coset=True, decimation=lib.DIT), ocpu.FFT_DIT_COSET),
                           ("FFTInverse(DIF, OnCoset)", dict(inverse=True, coset=True, decimation=lib.DIF), ocpu.FFT_INV_DIF_COSET)):
        y = x.clone()
        torch.cuda.synchronize()
        ctx.ntt_dev(y.data_ptr(), n, **kw)
        assert_limbs_equal(_host(y), ocpu.fft(xh, mode), f"{name} at 2^{logn}")


def test_transform_2p26_round_trip_and_oracle(ctx):
    """2^26, the top of config 3's sweep: coset DIT against oracle/c in full, and iNTT(DIF) o NTT(DIT) = id bit for bit
    (the two share no twiddle table: TW_INV vs TW_FWD)."""
    import torch
    logn = 26
    n = 1 << logn
    gen = torch.Generator(device="cuda:0")
    gen.manual_seed(3026)
    x = _rand_fr(torch, n, gen)
    # canonical inputs for the round trip: reduce below r by one multiplication with R mod r (Montgomery one)
    y = x.clone()
    torch.cuda.synchronize()
    ctx.ntt_dev(y.data_ptr(), n, inverse=False, coset=True, decimation=lib.DIT)
    assert_limbs_equal(_host(y), ocpu.fft(_host(x), ocpu.FFT_DIT_COSET), "FFT(DIT, OnCoset) at 2^26")
    z = y.clone()
    torch.cuda.synchronize()
    ctx.ntt_dev(z.data_ptr(), n, inverse=True, coset=False, decimation=lib.DIF)     # natural -> bit-reversed
    ctx.ntt_dev(z.data_ptr(), n, inverse=False, coset=False, decimation=lib.DIT)    # bit-reversed -> natural
    assert torch.equal(z, y), "FFT(DIT) o FFTInverse(DIF) is not the identity at 2^26"


@pytest.mark.parametrize("logn", [20, 24])
def test_compute_h_matches_oracle_at_large_sizes(ctx, logn):
    import torch
    n = 1 << logn
    nc = n - 3
    gen = torch.Generator(device="cuda:0")
    gen.manual_seed(3100 + logn)
    a, b = _rand_fr(torch, nc, gen), _rand_fr(torch, nc, gen)
    ah, bh = _host(a), _host(b)
    ch = ctx.fr_mul_vec(ah, bh)
    ch[::2] = _host(_rand_fr(torch, nc, gen))[::2]      # not divisible on half the rows: parity does not need it
    abc = torch.zeros((3, n, 4), dtype=torch.int64, device="cuda:0")
    for i, v in enumerate((ah, bh, ch)):
        abc[i, :nc] = torch.from_numpy(v.view(np.int64)).cuda()
    torch.cuda.synchronize()
    ctx.compute_h_dev(abc[0].data_ptr(), abc[1].data_ptr(), abc[2].data_ptr(), n)
    assert_limbs_equal(_host(abc[0]), ocpu.compute_h(ah, bh, ch, n), f"computeH at 2^{logn} vs oracle/c")


def _pow_table(x, logn):
    """[x^0, x^1, ..., x^(2^logn - 1)] (Montgomery limbs) by repeated doubling with the oracle's vector product"""
    pw = ocpu.fr_arr([1])
    xx = ocpu.fr_arr([x % R])
    for _ in range(logn):
        pw = np.concatenate([pw, ocpu.fr_mul_vec(pw, np.repeat(xx, len(pw), axis=0))])
        xx = ocpu.fr_mul_vec(xx, xx)
    return pw


def test_compute_h_2p26_closed_form(ctx):
    """2^26 (top of config 3): a, b are the evaluations of SPARSE polynomials A = a0 + a1 X^k, B = b0 + b1 X^m with
    k + m >= n, c = a o b.  Then A B - C = a1 b1 X^(k+m-n) (X^n - 1), so h must be the single coefficient a1 b1 at degree
    k + m - n -- in bit-reversed position -- and zero everywhere else.  O(n) oracle work, no oracle FFT."""
    import torch
    from oracle import ntt as ontt
    logn = 26
    n = 1 << logn
    w = ontt.Domain(n).gen                                     # the domain's generator (fft.Domain.Generator)
    assert pow(w, n, R) == 1 and pow(w, n // 2, R) != 1
    k, m = n - 12345, 3 * n // 4 + 77
    a0, a1, b0, b1 = 0x1111222233334444 % R, 0xABCDEF0123456789ABCDEF % R, 0x5555 % R, 0x987654321FEDCBA987654321 % R

    def term(c1, e):
        """c1 * X^e evaluated at w^i: (w^e)^i by repeated doubling, times c1 (oracle/c vector products)"""
        pw = _pow_table(pow(w, e, R), logn)
        return ocpu.fr_mul_vec(pw, np.repeat(ocpu.fr_arr([c1]), n, axis=0))

    av, bv = term(a1, k), term(b1, m)
    abc = torch.zeros((3, n, 4), dtype=torch.int64, device="cuda:0")
    a_full = _add_const(av, a0)
    b_full = _add_const(bv, b0)
    c_full = ocpu.fr_mul_vec(a_full, b_full)
    for i, v in enumerate((a_full, b_full, c_full)):
        abc[i] = torch.from_numpy(v.view(np.int64)).cuda()
    torch.cuda.synchronize()
    ctx.compute_h_dev(abc[0].data_ptr(), abc[1].data_ptr(), abc[2].data_ptr(), n)
    h = _host(abc[0])
    deg = k + m - n
    pos = int(bitrev_perm(n)[deg])
    want = np.zeros((n, 4), dtype=np.uint64)
    want[pos] = ocpu.fr_arr([a1 * b1 % R])[0]
    assert_limbs_equal(h, want, "computeH at 2^26 vs the closed-form quotient")


def _add_const(v, c):
    """v + c over Fr on Montgomery limb arrays, vectorised with numpy (c an int): add the limbs with carry, subtract r once"""
    cm = ocpu.fr_arr([c])[0]
    rl = np.array([(R >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)], dtype=np.uint64)
    out = np.empty_like(v)
    carry = np.zeros(len(v), dtype=np.uint64)
    for i in range(4):
        s = v[:, i] + cm[i]
        c1 = (s < v[:, i]).astype(np.uint64)
        s2 = s + carry
        c2 = (s2 < s).astype(np.uint64)
        out[:, i] = s2
        carry = c1 + c2
    # conditional subtraction of r (values < 2r)
    ge = np.ones(len(v), dtype=bool)
    decided = np.zeros(len(v), dtype=bool)
    for i in (3, 2, 1, 0):
        gt, lt = out[:, i] > rl[i], out[:, i] < rl[i]
        ge = np.where(~decided & lt, False, ge)
        decided |= gt | lt
    borrow = np.zeros(len(v), dtype=np.uint64)
    sub = np.empty_like(out)
    for i in range(4):
        d = out[:, i] - rl[i]
        b1 = (out[:, i] < rl[i]).astype(np.uint64)
        d2 = d - borrow
        b2 = (d < borrow).astype(np.uint64)
        sub[:, i] = d2
        borrow = b1 + b2
    out[ge] = sub[ge]
    return out


def test_prove_2p22_headline_config(ctx):
    """BASELINE config 4 itself: the MiMC-shaped chain at 2^22 constraints, key from the fixed-base kernel, all five
    points of g16_prove_bn254 against the trapdoor expectation with h from oracle/c's computeH."""
    nc = 1 << 22
    case = ChainCase(nc, seed=22, g1_mul=ctx.g1_fixed_base_mul, g2_mul=ctx.g2_fixed_base_mul)
    pk = case.pk
    handle = ctx.pk_upload(pk.domain_n, pk.g1_a, pk.g1_b, pk.g1_k, pk.g1_z, pk.g2_b, pk.infinity_a, pk.infinity_b, pk.nb_wires)
    try:
        pts, st = ctx.prove(handle, case.W, case.a, case.b, case.c)
    finally:
        ctx.pk_free(handle)
    h = ocpu.compute_h(case.a, case.b, case.c, case.n)
    assert_limbs_equal(pts, case.expected_points(h), "2^22 prove points vs trapdoor identity")
    assert st["msm_n"][0] == nc + 2 and st["msm_n"][3] == case.n - 1
