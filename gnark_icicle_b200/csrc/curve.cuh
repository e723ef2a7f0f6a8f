// Short-Weierstrass a=0 group arithmetic in XYZZ (extended Jacobian) coordinates, templated on the base
// field (Fp for G1, Fp2 for G2).  Used by the MSM kernels (device) and by the O(1) host epilogue.
//
// Replaces gnark-crypto G1Jac/G2Jac as used at backend/groth16/bn254/prove.go:204-304 and icicle's
// projective arithmetic behind iciclegnark.MsmOnDevice / MsmG2OnDevice
// (backend/groth16/bn254/icicle/icicle.go:302,315,332,355,382).  Only the affine result is an
// observable; XYZZ intermediates are private to this library.
// Formulas: EFD "xyzz" madd-2008-s (8M+2S), add-2008-s (12M+2S), dbl-2008-s-1, mdbl-2008-s-1.
#pragma once
#include "curve_select.h"
#include "field.cuh"

namespace g16 {

template <class BF>
struct Curve {
    typedef BF BaseField;
    typedef typename BF::T E;
    struct Aff { E x, y; };            // (0,0) = point at infinity (gnark layout)
    struct Xyzz { E x, y, zz, zzz; };  // zz == 0 = point at infinity

    G16_HD static bool is_inf(const Aff& p) { return BF::is_zero(p.x) && BF::is_zero(p.y); }
    G16_HD static bool is_inf(const Xyzz& p) { return BF::is_zero(p.zz); }
    G16_HD static Xyzz inf() { Xyzz r; r.x = BF::zero(); r.y = BF::zero(); r.zz = BF::zero(); r.zzz = BF::zero(); return r; }
    G16_HD static Aff aff_inf() { Aff r; r.x = BF::zero(); r.y = BF::zero(); return r; }
    G16_HD static Xyzz from_aff(const Aff& p) {
        if (is_inf(p)) return inf();
        Xyzz r; r.x = p.x; r.y = p.y; r.zz = BF::one(); r.zzz = BF::one(); return r;
    }
    G16_HD static Aff neg(const Aff& p) { Aff r; r.x = p.x; r.y = BF::neg(p.y); return r; }
    G16_HD static Xyzz neg(const Xyzz& p) { Xyzz r = p; r.y = BF::neg(p.y); return r; }

    // 2*P for affine P (mdbl-2008-s-1)
    G16_HD_COLD static Xyzz dbl_aff(const Aff& p) {
        if (is_inf(p)) return inf();
        E u = BF::dbl(p.y), v = BF::sqr(u), w = BF::mul(u, v), s = BF::mul(p.x, v);
        E x2 = BF::sqr(p.x), m = BF::add(BF::dbl(x2), x2);
        Xyzz r;
        r.x = BF::sub(BF::sub(BF::sqr(m), s), s);
        r.y = BF::sub(BF::mul(m, BF::sub(s, r.x)), BF::mul(w, p.y));
        r.zz = v; r.zzz = w;
        return r;
    }
    // 2*P (dbl-2008-s-1)
    G16_HD_COLD static Xyzz dbl(const Xyzz& p) {
        if (is_inf(p)) return p;
        E u = BF::dbl(p.y), v = BF::sqr(u), w = BF::mul(u, v), s = BF::mul(p.x, v);
        E x2 = BF::sqr(p.x), m = BF::add(BF::dbl(x2), x2);
        Xyzz r;
        r.x = BF::sub(BF::sub(BF::sqr(m), s), s);
        r.y = BF::sub(BF::mul(m, BF::sub(s, r.x)), BF::mul(w, p.y));
        r.zz = BF::mul(v, p.zz); r.zzz = BF::mul(w, p.zzz);
        return r;
    }
    // a += p, p affine (madd-2008-s); handles infinity, p == a (doubling) and p == -a
    G16_HD static void madd(Xyzz& a, const Aff& p) {
        if (is_inf(p)) return;
        if (is_inf(a)) { a.x = p.x; a.y = p.y; a.zz = BF::one(); a.zzz = BF::one(); return; }
        E pp = BF::sub(BF::mul(p.x, a.zz), a.x);
        E r = BF::sub(BF::mul(p.y, a.zzz), a.y);
        if (BF::is_zero(pp)) {
            if (BF::is_zero(r)) a = dbl_aff(p); else a = inf();
            return;
        }
        E p2 = BF::sqr(pp), p3 = BF::mul(pp, p2), q = BF::mul(a.x, p2);
        E x3 = BF::sub(BF::sub(BF::sub(BF::sqr(r), p3), q), q);
        a.y = BF::sub(BF::mul(r, BF::sub(q, x3)), BF::mul(a.y, p3));
        a.x = x3;
        a.zz = BF::mul(a.zz, p2);
        a.zzz = BF::mul(a.zzz, p3);
    }
    // a += b (add-2008-s)
    G16_HD_COLD static void add(Xyzz& a, const Xyzz& b) {
        if (is_inf(b)) return;
        if (is_inf(a)) { a = b; return; }
        E u1 = BF::mul(a.x, b.zz), u2 = BF::mul(b.x, a.zz);
        E s1 = BF::mul(a.y, b.zzz), s2 = BF::mul(b.y, a.zzz);
        E pp = BF::sub(u2, u1), r = BF::sub(s2, s1);
        if (BF::is_zero(pp)) {
            if (BF::is_zero(r)) a = dbl(a); else a = inf();
            return;
        }
        E p2 = BF::sqr(pp), p3 = BF::mul(pp, p2), q = BF::mul(u1, p2);
        E x3 = BF::sub(BF::sub(BF::sub(BF::sqr(r), p3), q), q);
        a.y = BF::sub(BF::mul(r, BF::sub(q, x3)), BF::mul(s1, p3));
        a.x = x3;
        a.zz = BF::mul(BF::mul(a.zz, b.zz), p2);
        a.zzz = BF::mul(BF::mul(a.zzz, b.zzz), p3);
    }
    // x = X/ZZ, y = Y/ZZZ with one inversion
    G16_HD_COLD static Aff to_aff(const Xyzz& p) {
        if (is_inf(p)) return aff_inf();
        E t = BF::inv(p.zzz);
        E zi = BF::mul(p.zz, t);  // = ZZ/ZZZ = 1/Z
        Aff r;
        r.x = BF::mul(p.x, BF::sqr(zi));
        r.y = BF::mul(p.y, t);
        return r;
    }
    // [k]P, k a plain little-endian integer (double-and-add; host epilogue and tiny device tails)
    G16_HD_COLD static Xyzz mul(const Xyzz& p, const uint32_t* k, int words) {
        Xyzz acc = inf();
        for (int i = words * 32 - 1; i >= 0; i--) {
            acc = dbl(acc);
            if ((k[i >> 5] >> (i & 31)) & 1) add(acc, p);
        }
        return acc;
    }
};

typedef Curve<FpC> G1;
typedef Curve<Fp2> G2;
// The same groups over the lazily reduced fields (coordinates in [0, 2p), field.cuh LazyFieldCall): used inside the
// bucket-accumulation loop only; Aff / Xyzz have the same memory layout as G1 / G2.
typedef Curve<FpL> G1L;
typedef Curve<Fp2L> G2L;
// MODE 0: every base-field multiplication is a call.  MODE 1: G1 inlines them (the loop body then has to stay a single
// group addition to fit the instruction cache), G2 calls at the Fp2 level.
template <class C, int MODE> struct LazyCurve;
template <> struct LazyCurve<G1, 0> { typedef G1L type; };
template <> struct LazyCurve<G2, 0> { typedef G2L type; };
template <> struct LazyCurve<G1, 1> { typedef Curve<FpLI> type; };
template <> struct LazyCurve<G2, 1> { typedef Curve<Fp2LC> type; };

}  // namespace g16
