/* Short-Weierstrass (a=0) group arithmetic, "templated" by macro over the base
 * field (oracle; TEST INFRASTRUCTURE ONLY).  Included twice by oracle_cpu.c:
 *   EC = g1, BF = fp   (E(Fp):  y^2 = x^3 + 3)
 *   EC = g2, BF = fp2  (E'(Fp2): y^2 = x^3 + 3/(9+u))
 * Restates the Jacobian formulas behind gnark-crypto's G1Jac/G2Jac as the
 * reference uses them (backend/groth16/bn254/prove.go:204-302: MultiExp,
 * AddMixed, AddAssign, ScalarMultiplication, FromJacobian).  Only the affine
 * result is comparable.
 */
#define CAT_(a, b) a##_##b
#define CAT(a, b) CAT_(a, b)
#define E(name) CAT(EC, name)
#define B(name) CAT(BF, name)
#define BT CAT(BF, t)

typedef struct { BT x, y; } E(aff);        /* (0,0) = infinity, gnark layout */
typedef struct { BT x, y, z; } E(jac);     /* z = 0 = infinity */

static inline int E(aff_is_inf)(const E(aff) *p) { return B(is_zero)(&p->x) && B(is_zero)(&p->y); }
static inline void E(jac_set_inf)(E(jac) *p) { B(set_one)(&p->x); B(set_one)(&p->y); B(set_zero)(&p->z); }
static inline int E(jac_is_inf)(const E(jac) *p) { return B(is_zero)(&p->z); }
static inline void E(from_aff)(E(jac) *r, const E(aff) *p) {
    if (E(aff_is_inf)(p)) { E(jac_set_inf)(r); return; }
    r->x = p->x; r->y = p->y; B(set_one)(&r->z);
}

static void E(jac_double)(E(jac) *r, const E(jac) *p) {
    if (E(jac_is_inf)(p)) { *r = *p; return; }
    BT A, Bq, C, D, Ee, F, t;
    B(sqr)(&A, &p->x); B(sqr)(&Bq, &p->y); B(sqr)(&C, &Bq);
    B(add)(&t, &p->x, &Bq); B(sqr)(&t, &t); B(sub)(&t, &t, &A); B(sub)(&t, &t, &C);
    B(add)(&D, &t, &t);
    B(add)(&Ee, &A, &A); B(add)(&Ee, &Ee, &A);
    B(sqr)(&F, &Ee);
    BT z3; B(mul)(&z3, &p->y, &p->z); B(add)(&z3, &z3, &z3);
    B(sub)(&r->x, &F, &D); B(sub)(&r->x, &r->x, &D);
    B(sub)(&t, &D, &r->x); B(mul)(&t, &t, &Ee);
    B(add)(&C, &C, &C); B(add)(&C, &C, &C); B(add)(&C, &C, &C);
    B(sub)(&r->y, &t, &C);
    r->z = z3;
}

static void E(jac_add)(E(jac) *r, const E(jac) *p, const E(jac) *q) {
    if (E(jac_is_inf)(p)) { *r = *q; return; }
    if (E(jac_is_inf)(q)) { *r = *p; return; }
    BT z1z1, z2z2, u1, u2, s1, s2, h, rr, hh, hhh, v, t;
    B(sqr)(&z1z1, &p->z); B(sqr)(&z2z2, &q->z);
    B(mul)(&u1, &p->x, &z2z2); B(mul)(&u2, &q->x, &z1z1);
    B(mul)(&s1, &p->y, &q->z); B(mul)(&s1, &s1, &z2z2);
    B(mul)(&s2, &q->y, &p->z); B(mul)(&s2, &s2, &z1z1);
    if (B(eq)(&u1, &u2)) {
        if (B(eq)(&s1, &s2)) { E(jac_double)(r, p); return; }
        E(jac_set_inf)(r); return;
    }
    B(sub)(&h, &u2, &u1); B(sub)(&rr, &s2, &s1);
    B(sqr)(&hh, &h); B(mul)(&hhh, &h, &hh); B(mul)(&v, &u1, &hh);
    BT x3, y3, z3;
    B(sqr)(&x3, &rr); B(sub)(&x3, &x3, &hhh); B(sub)(&x3, &x3, &v); B(sub)(&x3, &x3, &v);
    B(sub)(&t, &v, &x3); B(mul)(&y3, &rr, &t); B(mul)(&t, &s1, &hhh); B(sub)(&y3, &y3, &t);
    B(mul)(&z3, &p->z, &q->z); B(mul)(&z3, &z3, &h);
    r->x = x3; r->y = y3; r->z = z3;
}

static void E(jac_add_mixed)(E(jac) *r, const E(jac) *p, const E(aff) *q) {
    if (E(aff_is_inf)(q)) { *r = *p; return; }
    if (E(jac_is_inf)(p)) { E(from_aff)(r, q); return; }
    BT z1z1, u2, s2, h, rr, hh, hhh, v, t;
    B(sqr)(&z1z1, &p->z);
    B(mul)(&u2, &q->x, &z1z1);
    B(mul)(&s2, &q->y, &p->z); B(mul)(&s2, &s2, &z1z1);
    if (B(eq)(&p->x, &u2)) {
        if (B(eq)(&p->y, &s2)) { E(jac_double)(r, p); return; }
        E(jac_set_inf)(r); return;
    }
    B(sub)(&h, &u2, &p->x); B(sub)(&rr, &s2, &p->y);
    B(sqr)(&hh, &h); B(mul)(&hhh, &h, &hh); B(mul)(&v, &p->x, &hh);
    BT x3, y3, z3;
    B(sqr)(&x3, &rr); B(sub)(&x3, &x3, &hhh); B(sub)(&x3, &x3, &v); B(sub)(&x3, &x3, &v);
    B(sub)(&t, &v, &x3); B(mul)(&y3, &rr, &t); B(mul)(&t, &p->y, &hhh); B(sub)(&y3, &y3, &t);
    B(mul)(&z3, &p->z, &h);
    r->x = x3; r->y = y3; r->z = z3;
}

static void E(to_aff)(E(aff) *r, const E(jac) *p) {
    if (E(jac_is_inf)(p)) { B(set_zero)(&r->x); B(set_zero)(&r->y); return; }
    BT zi, zi2;
    B(inv)(&zi, &p->z); B(sqr)(&zi2, &zi);
    B(mul)(&r->x, &p->x, &zi2); B(mul)(&zi2, &zi2, &zi); B(mul)(&r->y, &p->y, &zi2);
}

/* r = [k] p, k = plain (non-Montgomery) 256-bit integer */
static void E(jac_mul)(E(jac) *r, const E(jac) *p, const fe *k) {
    E(jac) acc; E(jac_set_inf)(&acc);
    for (int i = 255; i >= 0; i--) {
        E(jac_double)(&acc, &acc);
        if ((k->l[i >> 6] >> (i & 63)) & 1) E(jac_add)(&acc, &acc, p);
    }
    *r = acc;
}

/* batch Jacobian -> affine with one inversion (Montgomery's trick) */
static void E(batch_to_aff)(E(aff) *out, const E(jac) *in, size_t n) {
    if (!n) return;
    BT *pref = (BT *)malloc(n * sizeof(BT));
    BT acc; B(set_one)(&acc);
    for (size_t i = 0; i < n; i++) {
        pref[i] = acc;
        if (!E(jac_is_inf)(&in[i])) B(mul)(&acc, &acc, &in[i].z);
    }
    BT ai; B(inv)(&ai, &acc);
    for (size_t i = n; i-- > 0;) {
        if (E(jac_is_inf)(&in[i])) { B(set_zero)(&out[i].x); B(set_zero)(&out[i].y); continue; }
        BT zi, zi2; B(mul)(&zi, &ai, &pref[i]); B(mul)(&ai, &ai, &in[i].z);
        B(sqr)(&zi2, &zi);
        B(mul)(&out[i].x, &in[i].x, &zi2); B(mul)(&zi2, &zi2, &zi); B(mul)(&out[i].y, &in[i].y, &zi2);
    }
    free(pref);
}

/* Pippenger bucket method over [lo,hi) of one c-bit window; scalars are plain integers.
 * This is the algorithm family of gnark-crypto's MultiExp (bucket method, per-window tasks). */
static void E(msm_window)(E(jac) *out, const E(aff) *pts, const fe *sc, size_t n, int c, int w) {
    size_t nb = ((size_t)1 << c) - 1;
    E(jac) *bk = (E(jac) *)malloc(nb * sizeof(E(jac)));
    for (size_t i = 0; i < nb; i++) E(jac_set_inf)(&bk[i]);
    int bit = w * c;
    for (size_t i = 0; i < n; i++) {
        int limb = bit >> 6, sh = bit & 63;
        uint64_t d = limb < 4 ? sc[i].l[limb] >> sh : 0;
        if (sh && limb + 1 < 4 && sh + c > 64) d |= sc[i].l[limb + 1] << (64 - sh);
        d &= nb;
        if (d) E(jac_add_mixed)(&bk[d - 1], &bk[d - 1], &pts[i]);
    }
    E(jac) run, acc; E(jac_set_inf)(&run); E(jac_set_inf)(&acc);
    for (size_t i = nb; i-- > 0;) { E(jac_add)(&run, &run, &bk[i]); E(jac_add)(&acc, &acc, &run); }
    free(bk);
    *out = acc;
}

#undef E
#undef B
#undef BT
#undef CAT
#undef CAT_
