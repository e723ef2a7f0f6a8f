/* Montgomery prime field on NL 64-bit limbs, "templated" by macro (oracle; TEST INFRASTRUCTURE ONLY).
 * Included by field.h once per limb count with
 *   FE   element type        FT   field-description type     NL  limbs
 *   FEN(x) / RAW(x) / FN(x)  names of the element, raw-integer and field functions
 * Restates gnark-crypto's fr/fp element arithmetic (CIOS Montgomery product, conditional subtraction);
 * see field.h for the provenance note.
 */
typedef struct { uint64_t l[NL]; } FE;                   /* one field element, little-endian limbs */
typedef struct { FE m; uint64_t inv; FE r2; FE one; } FT; /* modulus, -m^-1 mod 2^64, R^2 mod m, R mod m */

static inline int FEN(is_zero)(const FE *a) {
    uint64_t o = 0;
    for (int i = 0; i < NL; i++) o |= a->l[i];
    return o == 0;
}
static inline int FEN(eq)(const FE *a, const FE *b) { return memcmp(a, b, sizeof(FE)) == 0; }
static inline int FEN(geq)(const FE *a, const FE *b) {
    for (int i = NL - 1; i >= 0; i--) {
        if (a->l[i] > b->l[i]) return 1;
        if (a->l[i] < b->l[i]) return 0;
    }
    return 1;
}
static inline uint64_t RAW(add)(FE *r, const FE *a, const FE *b) {
    u128 c = 0;
    for (int i = 0; i < NL; i++) { c += (u128)a->l[i] + b->l[i]; r->l[i] = (uint64_t)c; c >>= 64; }
    return (uint64_t)c;
}
static inline uint64_t RAW(sub)(FE *r, const FE *a, const FE *b) {
    uint64_t br = 0;
    for (int i = 0; i < NL; i++) {
        u128 d = (u128)a->l[i] - b->l[i] - br;
        r->l[i] = (uint64_t)d; br = (uint64_t)(d >> 64) & 1;
    }
    return br;
}
static inline void FN(add)(const FT *F, FE *r, const FE *a, const FE *b) {
    FE t; uint64_t c = RAW(add)(&t, a, b);
    if (c || FEN(geq)(&t, &F->m)) RAW(sub)(&t, &t, &F->m);
    *r = t;
}
static inline void FN(sub)(const FT *F, FE *r, const FE *a, const FE *b) {
    FE t; if (RAW(sub)(&t, a, b)) RAW(add)(&t, &t, &F->m);
    *r = t;
}
static inline void FN(neg)(const FT *F, FE *r, const FE *a) {
    if (FEN(is_zero)(a)) { *r = *a; return; }
    RAW(sub)(r, &F->m, a);
}
static inline void FN(dbl)(const FT *F, FE *r, const FE *a) { FN(add)(F, r, a, a); }

/* CIOS Montgomery multiplication */
static inline void FN(mul)(const FT *F, FE *r, const FE *a, const FE *b) {
    uint64_t t[NL + 2];
    memset(t, 0, sizeof t);
    for (int i = 0; i < NL; i++) {
        u128 c = 0;
        for (int j = 0; j < NL; j++) {
            c += (u128)a->l[j] * b->l[i] + t[j];
            t[j] = (uint64_t)c; c >>= 64;
        }
        c += t[NL]; t[NL] = (uint64_t)c; t[NL + 1] = (uint64_t)(c >> 64);
        uint64_t m = t[0] * F->inv;
        c = (u128)m * F->m.l[0] + t[0]; c >>= 64;
        for (int j = 1; j < NL; j++) {
            c += (u128)m * F->m.l[j] + t[j];
            t[j - 1] = (uint64_t)c; c >>= 64;
        }
        c += t[NL]; t[NL - 1] = (uint64_t)c; t[NL] = t[NL + 1] + (uint64_t)(c >> 64);
    }
    FE o; memcpy(&o, t, sizeof o);
    if (t[NL] || FEN(geq)(&o, &F->m)) RAW(sub)(&o, &o, &F->m);
    *r = o;
}
static inline void FN(sqr)(const FT *F, FE *r, const FE *a) { FN(mul)(F, r, a, a); }
static inline void FN(to_mont)(const FT *F, FE *r, const FE *a) { FN(mul)(F, r, a, &F->r2); }
static inline void FN(from_mont)(const FT *F, FE *r, const FE *a) {
    FE one; memset(&one, 0, sizeof one); one.l[0] = 1; FN(mul)(F, r, a, &one);
}
/* r = a^e, e plain NL-limb integer, a Montgomery */
static inline void FN(pow)(const FT *F, FE *r, const FE *a, const FE *e) {
    FE acc = F->one, base = *a;
    for (int i = 0; i < 64 * NL; i++) {
        if ((e->l[i >> 6] >> (i & 63)) & 1) FN(mul)(F, &acc, &acc, &base);
        FN(sqr)(F, &base, &base);
    }
    *r = acc;
}
static inline void FN(inv)(const FT *F, FE *r, const FE *a) {
    FE e = F->m, two; memset(&two, 0, sizeof two); two.l[0] = 2;
    RAW(sub)(&e, &e, &two);
    FN(pow)(F, r, a, &e);
}
static inline void FN(set_u64)(const FT *F, FE *r, uint64_t v) {
    FE t; memset(&t, 0, sizeof t); t.l[0] = v; FN(to_mont)(F, r, &t);
}
