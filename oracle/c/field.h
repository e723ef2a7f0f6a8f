/* 256-bit Montgomery fields on 4x64 limbs (oracle; TEST INFRASTRUCTURE ONLY).
 *
 * CPU restatement of the arithmetic the reference obtains from gnark-crypto
 * (github.com/consensys/gnark-crypto v0.14.1-0.20241010154951-6638408a49f3,
 * not vendored; call sites backend/groth16/bn254/prove.go:356-399 for fr,
 * :202-304 for curve ops).  Element layout = gnark-crypto's: 4 little-endian
 * uint64 limbs, Montgomery form with R = 2^256.  Parity status: unpinned by
 * golden vectors (see oracle/__init__.py); validated against the Python oracle.
 */
#ifndef ORC_FIELD_H
#define ORC_FIELD_H
#include <stdint.h>
#include <string.h>

typedef unsigned __int128 u128;
typedef struct { uint64_t l[4]; } fe;          /* one field element */
typedef struct { fe m; uint64_t inv; fe r2; fe one; } field_t;

/* BN254 moduli, -m^-1 mod 2^64, R^2 mod m, R mod m */
static const field_t FR = {
    {{0x43e1f593f0000001ULL, 0x2833e84879b97091ULL, 0xb85045b68181585dULL, 0x30644e72e131a029ULL}},
    0xc2e1f593efffffffULL,
    {{0x1bb8e645ae216da7ULL, 0x53fe3ab1e35c59e3ULL, 0x8c49833d53bb8085ULL, 0x0216d0b17f4e44a5ULL}},
    {{0xac96341c4ffffffbULL, 0x36fc76959f60cd29ULL, 0x666ea36f7879462eULL, 0x0e0a77c19a07df2fULL}}};
static const field_t FP = {
    {{0x3c208c16d87cfd47ULL, 0x97816a916871ca8dULL, 0xb85045b68181585dULL, 0x30644e72e131a029ULL}},
    0x87d20782e4866389ULL,
    {{0xf32cfc5b538afa89ULL, 0xb5e71911d44501fbULL, 0x47ab1eff0a417ff6ULL, 0x06d89f71cab8351fULL}},
    {{0xd35d438dc58f0d9dULL, 0x0a78eb28f5c70b3dULL, 0x666ea36f7879462cULL, 0x0e0a77c19a07df2fULL}}};

static inline int fe_is_zero(const fe *a) { return (a->l[0] | a->l[1] | a->l[2] | a->l[3]) == 0; }
static inline int fe_eq(const fe *a, const fe *b) { return memcmp(a, b, sizeof(fe)) == 0; }

static inline int fe_geq(const fe *a, const fe *b) {
    for (int i = 3; i >= 0; i--) {
        if (a->l[i] > b->l[i]) return 1;
        if (a->l[i] < b->l[i]) return 0;
    }
    return 1;
}
static inline uint64_t raw_add(fe *r, const fe *a, const fe *b) {
    u128 c = 0;
    for (int i = 0; i < 4; i++) { c += (u128)a->l[i] + b->l[i]; r->l[i] = (uint64_t)c; c >>= 64; }
    return (uint64_t)c;
}
static inline uint64_t raw_sub(fe *r, const fe *a, const fe *b) {
    uint64_t br = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)a->l[i] - b->l[i] - br;
        r->l[i] = (uint64_t)d; br = (uint64_t)(d >> 64) & 1;
    }
    return br;
}
static inline void f_add(const field_t *F, fe *r, const fe *a, const fe *b) {
    fe t; uint64_t c = raw_add(&t, a, b);
    if (c || fe_geq(&t, &F->m)) raw_sub(&t, &t, &F->m);
    *r = t;
}
static inline void f_sub(const field_t *F, fe *r, const fe *a, const fe *b) {
    fe t; if (raw_sub(&t, a, b)) raw_add(&t, &t, &F->m);
    *r = t;
}
static inline void f_neg(const field_t *F, fe *r, const fe *a) {
    if (fe_is_zero(a)) { *r = *a; return; }
    raw_sub(r, &F->m, a);
}
static inline void f_dbl(const field_t *F, fe *r, const fe *a) { f_add(F, r, a, a); }

/* CIOS Montgomery multiplication */
static inline void f_mul(const field_t *F, fe *r, const fe *a, const fe *b) {
    uint64_t t[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 4; i++) {
        u128 c = 0;
        for (int j = 0; j < 4; j++) {
            c += (u128)a->l[j] * b->l[i] + t[j];
            t[j] = (uint64_t)c; c >>= 64;
        }
        c += t[4]; t[4] = (uint64_t)c; t[5] = (uint64_t)(c >> 64);
        uint64_t m = t[0] * F->inv;
        c = (u128)m * F->m.l[0] + t[0]; c >>= 64;
        for (int j = 1; j < 4; j++) {
            c += (u128)m * F->m.l[j] + t[j];
            t[j - 1] = (uint64_t)c; c >>= 64;
        }
        c += t[4]; t[3] = (uint64_t)c; t[4] = t[5] + (uint64_t)(c >> 64);
    }
    fe o = {{t[0], t[1], t[2], t[3]}};
    if (t[4] || fe_geq(&o, &F->m)) raw_sub(&o, &o, &F->m);
    *r = o;
}
static inline void f_sqr(const field_t *F, fe *r, const fe *a) { f_mul(F, r, a, a); }
static inline void f_to_mont(const field_t *F, fe *r, const fe *a) { f_mul(F, r, a, &F->r2); }
static inline void f_from_mont(const field_t *F, fe *r, const fe *a) {
    fe one = {{1, 0, 0, 0}}; f_mul(F, r, a, &one);
}
/* r = a^e, e plain 4-limb integer, a Montgomery */
static inline void f_pow(const field_t *F, fe *r, const fe *a, const fe *e) {
    fe acc = F->one, base = *a;
    for (int i = 0; i < 256; i++) {
        if ((e->l[i >> 6] >> (i & 63)) & 1) f_mul(F, &acc, &acc, &base);
        f_sqr(F, &base, &base);
    }
    *r = acc;
}
static inline void f_inv(const field_t *F, fe *r, const fe *a) {
    fe e = F->m, two = {{2, 0, 0, 0}};
    raw_sub(&e, &e, &two);
    f_pow(F, r, a, &e);
}
static inline void f_set_u64(const field_t *F, fe *r, uint64_t v) {
    fe t = {{v, 0, 0, 0}}; f_to_mont(F, r, &t);
}
#endif
