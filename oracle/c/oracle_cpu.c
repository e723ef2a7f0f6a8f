/* Multithreaded CPU oracle for the Groth16 BN254 hot path.
 *
 * (ORC_CURVE=377 builds the same file for BLS12-377, BASELINE.json config 5:
 *  backend/groth16/bls12-377/prove.go is bn254/prove.go over the other curve package.)
 *
 * TEST INFRASTRUCTURE ONLY -- never linked into or called from the product
 * library; used by tests/ as the checker at sizes Python cannot reach and by
 * bench.py as the timed CPU baseline ("port": a restatement of the
 * gnark-crypto algorithms, NOT gnark-crypto, which needs Go and x86 assembly).
 *
 * Restates, from the reference's call sites:
 *   computeH            backend/groth16/bn254/prove.go:356-399
 *   G1/G2 MultiExp      backend/groth16/bn254/prove.go:204,217,237,247,293
 *   BatchScalarMultiplicationG1/G2   setup.go:251,320
 *   setupABC / Z / K scalars         setup.go:143-210,364-446
 * Parity status: unpinned by golden vectors (none exist in the reference for
 * this path); validated against the Python oracle (tests/test_oracle_c.py).
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>
#include "field.h"

/* ------------------------------------------------------------ fp / fp2 adapters */
#if ORC_CURVE == 254
typedef fe fp_t;
#define FPF(x) f_##x
#define FPE(x) fe_##x
#define FP2_NONRES 1    /* Fp2 = Fp[u]/(u^2 + 1) */
#define G1_B 3          /* y^2 = x^3 + 3 */
#define FR_COSET_GEN 5  /* fft.Domain FrMultiplicativeGen */
#define FR_ADICITY 28
#else
typedef fe6 fp_t;
#define FPF(x) f6_##x
#define FPE(x) fe6_##x
#define FP2_NONRES 5    /* Fp2 = Fp[u]/(u^2 + 5) (gnark-crypto bls12-377 E2) */
#define G1_B 1          /* y^2 = x^3 + 1 */
#define FR_COSET_GEN 22
#define FR_ADICITY 47
#endif
static inline void fp_add(fp_t *r, const fp_t *a, const fp_t *b) { FPF(add)(&FP, r, a, b); }
static inline void fp_sub(fp_t *r, const fp_t *a, const fp_t *b) { FPF(sub)(&FP, r, a, b); }
static inline void fp_mul(fp_t *r, const fp_t *a, const fp_t *b) { FPF(mul)(&FP, r, a, b); }
static inline void fp_sqr(fp_t *r, const fp_t *a) { FPF(mul)(&FP, r, a, a); }
static inline void fp_inv(fp_t *r, const fp_t *a) { FPF(inv)(&FP, r, a); }
static inline void fp_neg(fp_t *r, const fp_t *a) { FPF(neg)(&FP, r, a); }
static inline int fp_is_zero(const fp_t *a) { return FPE(is_zero)(a); }
static inline int fp_eq(const fp_t *a, const fp_t *b) { return FPE(eq)(a, b); }
static inline void fp_set_one(fp_t *r) { *r = FP.one; }
static inline void fp_set_zero(fp_t *r) { memset(r, 0, sizeof *r); }
/* r = k*a for the small non-residue */
static inline void fp_mul_nonres(fp_t *r, const fp_t *a) {
#if FP2_NONRES == 1
    *r = *a;
#else
    fp_t t, u; fp_add(&t, a, a); fp_add(&t, &t, &t); fp_add(&u, &t, a); *r = u;   /* 5a */
#endif
}

typedef struct { fp_t a0, a1; } fp2_t;   /* a0 + a1*u, u^2 = -FP2_NONRES (gnark E2{A0,A1}) */
static inline void fp2_add(fp2_t *r, const fp2_t *a, const fp2_t *b) { fp_add(&r->a0, &a->a0, &b->a0); fp_add(&r->a1, &a->a1, &b->a1); }
static inline void fp2_sub(fp2_t *r, const fp2_t *a, const fp2_t *b) { fp_sub(&r->a0, &a->a0, &b->a0); fp_sub(&r->a1, &a->a1, &b->a1); }
static inline void fp2_mul(fp2_t *r, const fp2_t *a, const fp2_t *b) {
    fp_t t0, t1, t2, t3;
    fp_mul(&t0, &a->a0, &b->a0); fp_mul(&t1, &a->a1, &b->a1);
    fp_mul(&t2, &a->a0, &b->a1); fp_mul(&t3, &a->a1, &b->a0);
    fp_mul_nonres(&t1, &t1);
    fp_sub(&r->a0, &t0, &t1); fp_add(&r->a1, &t2, &t3);
}
static inline void fp2_sqr(fp2_t *r, const fp2_t *a) { fp2_mul(r, a, a); }
static inline void fp2_inv(fp2_t *r, const fp2_t *a) {      /* conj(a) / (a0^2 + NONRES a1^2) */
    fp_t t0, t1; fp_sqr(&t0, &a->a0); fp_sqr(&t1, &a->a1); fp_mul_nonres(&t1, &t1); fp_add(&t0, &t0, &t1); fp_inv(&t0, &t0);
    fp_mul(&r->a0, &a->a0, &t0); fp_mul(&t1, &a->a1, &t0); fp_neg(&r->a1, &t1);
}
static inline int fp2_is_zero(const fp2_t *a) { return fp_is_zero(&a->a0) && fp_is_zero(&a->a1); }
static inline int fp2_eq(const fp2_t *a, const fp2_t *b) { return fp_eq(&a->a0, &b->a0) && fp_eq(&a->a1, &b->a1); }
static inline void fp2_set_one(fp2_t *r) { r->a0 = FP.one; memset(&r->a1, 0, sizeof(fp_t)); }
static inline void fp2_set_zero(fp2_t *r) { memset(r, 0, sizeof *r); }

#define EC g1
#define BF fp
#include "ec_tmpl.h"
#undef EC
#undef BF
#define EC g2
#define BF fp2
#include "ec_tmpl.h"
#undef EC
#undef BF

/* ------------------------------------------------------------ threading */
typedef void (*range_fn)(size_t lo, size_t hi, int tid, void *ctx);
typedef struct { range_fn fn; void *ctx; size_t lo, hi; int tid; } job_t;
static void *job_tramp(void *p) { job_t *j = (job_t *)p; j->fn(j->lo, j->hi, j->tid, j->ctx); return NULL; }

int orc_threads(void) { long n = sysconf(_SC_NPROCESSORS_ONLN); return n > 0 ? (int)n : 1; }

/* internal/utils/parallelize.go:9-44: split [0,n) into nthreads contiguous chunks */
static void par_for(size_t n, int nthreads, range_fn fn, void *ctx) {
    if (nthreads < 1) nthreads = 1;
    if ((size_t)nthreads > n) nthreads = n ? (int)n : 1;
    if (nthreads == 1) { fn(0, n, 0, ctx); return; }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * nthreads);
    job_t *jobs = (job_t *)malloc(sizeof(job_t) * nthreads);
    size_t per = n / nthreads, extra = n % nthreads, pos = 0;
    for (int t = 0; t < nthreads; t++) {
        size_t len = per + ((size_t)t < extra ? 1 : 0);
        jobs[t] = (job_t){fn, ctx, pos, pos + len, t}; pos += len;
        pthread_create(&th[t], NULL, job_tramp, &jobs[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    free(th); free(jobs);
}

/* ------------------------------------------------------------ PRNG (oracle/prng.py) */
typedef struct { uint64_t s[4]; } xo_t;
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static void xo_seed(xo_t *g, uint64_t seed) {
    for (int i = 0; i < 4; i++) {
        seed += 0x9E3779B97F4A7C15ULL;
        uint64_t z = seed;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        g->s[i] = z ^ (z >> 31);
    }
}
static inline uint64_t xo_next(xo_t *g) {
    uint64_t *s = g->s, result = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
    return result;
}
/* 4 draws -> 256-bit integer mod m -> Montgomery */
static void xo_field(xo_t *g, const field_t *F, fe *out) {
    fe v; for (int i = 0; i < 4; i++) v.l[i] = xo_next(g);
    while (fe_geq(&v, &F->m)) raw_sub(&v, &v, &F->m);
    f_to_mont(F, out, &v);
}
void orc_fr_random(uint64_t seed, uint64_t n, fe *out) {
    xo_t g; xo_seed(&g, seed);
    for (uint64_t i = 0; i < n; i++) xo_field(&g, &FR, &out[i]);
}

/* ------------------------------------------------------------ Fr vector helpers */
void orc_fr_mul_vec(const fe *a, const fe *b, uint64_t n, fe *out) { for (uint64_t i = 0; i < n; i++) f_mul(&FR, &out[i], &a[i], &b[i]); }
void orc_fr_dot(const fe *a, const fe *b, uint64_t n, fe *out) {
    fe acc; memset(&acc, 0, sizeof acc);
    for (uint64_t i = 0; i < n; i++) { fe t; f_mul(&FR, &t, &a[i], &b[i]); f_add(&FR, &acc, &acc, &t); }
    *out = acc;
}

/* ------------------------------------------------------------ generators */
#if ORC_CURVE == 254
/* G1 = (1, 2); G2 = the EIP-197 / gnark-crypto bn254.Generators point (plain integers) */
static const fp_t G1X = {{1, 0, 0, 0}}, G1Y = {{2, 0, 0, 0}};
static const fp_t G2X0 = {{0x46debd5cd992f6edULL, 0x674322d4f75edaddULL, 0x426a00665e5c4479ULL, 0x1800deef121f1e76ULL}};
static const fp_t G2X1 = {{0x97e485b7aef312c2ULL, 0xf1aa493335a9e712ULL, 0x7260bfb731fb5d25ULL, 0x198e9393920d483aULL}};
static const fp_t G2Y0 = {{0x4ce6cc0166fa7daaULL, 0xe3d1e7690c43d37bULL, 0x4aab71808dcb408fULL, 0x12c85ea5db8c6debULL}};
static const fp_t G2Y1 = {{0x55acdadcd122975bULL, 0xbc4b313370b38ef3ULL, 0xec9e99ad690c3395ULL, 0x090689d0585ff075ULL}};
#else
/* gnark-crypto's bls12-377 G1 generator; the G2 point is the harness generator of oracle/curve.py (cofactor times
 * the twist point with x = 2): every G2 operand of the proving path comes from the key, so only the harness's own
 * keys depend on it.  Plain integers. */
static const fp_t G1X = {{0xeab9b16eb21be9efULL, 0xd5481512ffcd394eULL, 0x188282c8bd37cb5cULL, 0x85951e2caa9d41bbULL, 0xc8fc6225bf87ff54ULL, 0x008848defe740a67ULL}};
static const fp_t G1Y = {{0xfd82de55559c8ea6ULL, 0xc2fe3d3634a9591aULL, 0x6d182ad44fb82305ULL, 0xbd7fb348ca3e52d9ULL, 0x1f674f5d30afeec4ULL, 0x01914a69c5102effULL}};
static const fp_t G2X0 = {{0x7faab58c9b87bda9ULL, 0x20a7c4c78130d2d0ULL, 0x401724764cd34f4dULL, 0x049f123bf27dce03ULL, 0x35bb08884aee7b99ULL, 0x00415abb1137c889ULL}};
static const fp_t G2X1 = {{0x0ddf884b60d27f88ULL, 0x4de60b0c0f56b5e1ULL, 0xa208c4859cf3630fULL, 0x24148cd0d42beba1ULL, 0x9d9f07b86366d6eaULL, 0x00a272c24ca7d342ULL}};
static const fp_t G2Y0 = {{0x76521ebbbb23ee23ULL, 0x33cc6492b66b5ee5ULL, 0x4ac4ef1a54a524d9ULL, 0xf523908ed09da02fULL, 0xb97bc0e7237bbb97ULL, 0x0199275defaacc0dULL}};
static const fp_t G2Y1 = {{0xeb2f80ac9c3a8051ULL, 0xe66e6d2df31cbe0cULL, 0x78536c115ac38999ULL, 0x509504b67ef09df5ULL, 0x4752e212f85ed167ULL, 0x0121cb2668f07eaaULL}};
#endif

static void g1_gen(g1_aff *g) { FPF(to_mont)(&FP, &g->x, &G1X); FPF(to_mont)(&FP, &g->y, &G1Y); }
static void g2_gen(g2_aff *g) {
    FPF(to_mont)(&FP, &g->x.a0, &G2X0); FPF(to_mont)(&FP, &g->x.a1, &G2X1);
    FPF(to_mont)(&FP, &g->y.a0, &G2Y0); FPF(to_mont)(&FP, &g->y.a1, &G2Y1);
}

/* ------------------------------------------------------------ fixed-base [k]G, 8-bit windows
 * (curve.BatchScalarMultiplicationG1/G2, setup.go:251,320) */
#define FB_W 8
#define FB_NW 32
#define FB_ROW 255
static g1_aff *fb1; static g2_aff *fb2;
static pthread_mutex_t fb_lock = PTHREAD_MUTEX_INITIALIZER;

static void fb1_init(void) {
    pthread_mutex_lock(&fb_lock);
    if (!fb1) {
        g1_jac *t = (g1_jac *)malloc(sizeof(g1_jac) * FB_NW * FB_ROW);
        g1_aff g; g1_gen(&g); g1_jac base; g1_from_aff(&base, &g);
        for (int j = 0; j < FB_NW; j++) {
            t[j * FB_ROW] = base;
            for (int d = 1; d < FB_ROW; d++) g1_jac_add(&t[j * FB_ROW + d], &t[j * FB_ROW + d - 1], &base);
            g1_jac nb; g1_jac_add(&nb, &t[j * FB_ROW + FB_ROW - 1], &base); base = nb;
        }
        g1_aff *a = (g1_aff *)malloc(sizeof(g1_aff) * FB_NW * FB_ROW);
        g1_batch_to_aff(a, t, FB_NW * FB_ROW); free(t); fb1 = a;
    }
    pthread_mutex_unlock(&fb_lock);
}
static void fb2_init(void) {
    pthread_mutex_lock(&fb_lock);
    if (!fb2) {
        g2_jac *t = (g2_jac *)malloc(sizeof(g2_jac) * FB_NW * FB_ROW);
        g2_aff g; g2_gen(&g); g2_jac base; g2_from_aff(&base, &g);
        for (int j = 0; j < FB_NW; j++) {
            t[j * FB_ROW] = base;
            for (int d = 1; d < FB_ROW; d++) g2_jac_add(&t[j * FB_ROW + d], &t[j * FB_ROW + d - 1], &base);
            g2_jac nb; g2_jac_add(&nb, &t[j * FB_ROW + FB_ROW - 1], &base); base = nb;
        }
        g2_aff *a = (g2_aff *)malloc(sizeof(g2_aff) * FB_NW * FB_ROW);
        g2_batch_to_aff(a, t, FB_NW * FB_ROW); free(t); fb2 = a;
    }
    pthread_mutex_unlock(&fb_lock);
}

typedef struct { const fe *sc; void *out; } fb_ctx;
#define FB_CHUNK 1024
static void fb1_range(size_t lo, size_t hi, int tid, void *vc) {
    (void)tid; fb_ctx *c = (fb_ctx *)vc; g1_aff *out = (g1_aff *)c->out;
    g1_jac *buf = (g1_jac *)malloc(sizeof(g1_jac) * FB_CHUNK);
    for (size_t s = lo; s < hi; s += FB_CHUNK) {
        size_t e = s + FB_CHUNK < hi ? s + FB_CHUNK : hi;
        for (size_t i = s; i < e; i++) {
            fe k; f_from_mont(&FR, &k, &c->sc[i]);
            g1_jac acc; g1_jac_set_inf(&acc);
            for (int j = 0; j < FB_NW; j++) {
                unsigned d = (k.l[j >> 3] >> ((j & 7) * 8)) & 0xff;
                if (d) g1_jac_add_mixed(&acc, &acc, &fb1[j * FB_ROW + d - 1]);
            }
            buf[i - s] = acc;
        }
        g1_batch_to_aff(out + s, buf, e - s);
    }
    free(buf);
}
static void fb2_range(size_t lo, size_t hi, int tid, void *vc) {
    (void)tid; fb_ctx *c = (fb_ctx *)vc; g2_aff *out = (g2_aff *)c->out;
    g2_jac *buf = (g2_jac *)malloc(sizeof(g2_jac) * FB_CHUNK);
    for (size_t s = lo; s < hi; s += FB_CHUNK) {
        size_t e = s + FB_CHUNK < hi ? s + FB_CHUNK : hi;
        for (size_t i = s; i < e; i++) {
            fe k; f_from_mont(&FR, &k, &c->sc[i]);
            g2_jac acc; g2_jac_set_inf(&acc);
            for (int j = 0; j < FB_NW; j++) {
                unsigned d = (k.l[j >> 3] >> ((j & 7) * 8)) & 0xff;
                if (d) g2_jac_add_mixed(&acc, &acc, &fb2[j * FB_ROW + d - 1]);
            }
            buf[i - s] = acc;
        }
        g2_batch_to_aff(out + s, buf, e - s);
    }
    free(buf);
}
/* out[i] = [scalars[i]] G  (scalars Montgomery Fr; out affine Montgomery) */
void orc_g1_mul_gen_batch(const fe *scalars, uint64_t n, g1_aff *out, int threads) {
    fb1_init(); fb_ctx c = {scalars, out}; par_for(n, threads, fb1_range, &c);
}
void orc_g2_mul_gen_batch(const fe *scalars, uint64_t n, g2_aff *out, int threads) {
    fb2_init(); fb_ctx c = {scalars, out}; par_for(n, threads, fb2_range, &c);
}

/* ------------------------------------------------------------ MSM (bucket method)
 * gnark-crypto MultiExp family: c-bit windows, one task per (window, point chunk),
 * running-sum bucket reduction, Horner combination of the windows. */
static int msm_pick_c(uint64_t n) {
    int c = 4; while (((uint64_t)1 << (c + 4)) < n && c < 16) c++;   /* ~log2(n)-4, clamped */
    return c;
}
typedef struct { const void *pts; const fe *sc; size_t n; int c, nwin, nchunk; void *partial; } msm_ctx;
static void msm1_task(size_t lo, size_t hi, int tid, void *vc) {
    (void)tid; msm_ctx *m = (msm_ctx *)vc;
    for (size_t t = lo; t < hi; t++) {
        int w = (int)(t / m->nchunk), ch = (int)(t % m->nchunk);
        size_t a = m->n * ch / m->nchunk, b = m->n * (ch + 1) / m->nchunk;
        g1_msm_window(&((g1_jac *)m->partial)[t], (const g1_aff *)m->pts + a, m->sc + a, b - a, m->c, w);
    }
}
static void msm2_task(size_t lo, size_t hi, int tid, void *vc) {
    (void)tid; msm_ctx *m = (msm_ctx *)vc;
    for (size_t t = lo; t < hi; t++) {
        int w = (int)(t / m->nchunk), ch = (int)(t % m->nchunk);
        size_t a = m->n * ch / m->nchunk, b = m->n * (ch + 1) / m->nchunk;
        g2_msm_window(&((g2_jac *)m->partial)[t], (const g2_aff *)m->pts + a, m->sc + a, b - a, m->c, w);
    }
}
typedef struct { const fe *in; fe *out; } conv_ctx;
static void from_mont_range(size_t lo, size_t hi, int tid, void *vc) {
    (void)tid; conv_ctx *c = (conv_ctx *)vc;
    for (size_t i = lo; i < hi; i++) f_from_mont(&FR, &c->out[i], &c->in[i]);
}
/* out = sum scalars[i]*points[i]; scalars Montgomery Fr, points/out affine Montgomery */
void orc_msm_g1(const g1_aff *pts, const fe *scalars, uint64_t n, g1_aff *out, int c, int threads) {
    if (!n) { memset(out, 0, sizeof *out); return; }
    if (c <= 0) c = msm_pick_c(n);
    fe *sc = (fe *)malloc(sizeof(fe) * n);
    conv_ctx cv = {scalars, sc}; par_for(n, threads, from_mont_range, &cv);
    int nwin = (254 + c - 1) / c, nchunk = 1;
    while (nwin * nchunk < threads && (n / (nchunk + 1)) > ((uint64_t)4 << c)) nchunk++;
    g1_jac *partial = (g1_jac *)malloc(sizeof(g1_jac) * nwin * nchunk);
    msm_ctx m = {pts, sc, n, c, nwin, nchunk, partial};
    par_for((size_t)nwin * nchunk, threads, msm1_task, &m);
    g1_jac acc; g1_jac_set_inf(&acc);
    for (int w = nwin - 1; w >= 0; w--) {
        for (int i = 0; i < c; i++) g1_jac_double(&acc, &acc);
        for (int ch = 0; ch < nchunk; ch++) g1_jac_add(&acc, &acc, &partial[w * nchunk + ch]);
    }
    g1_to_aff(out, &acc); free(partial); free(sc);
}
void orc_msm_g2(const g2_aff *pts, const fe *scalars, uint64_t n, g2_aff *out, int c, int threads) {
    if (!n) { memset(out, 0, sizeof *out); return; }
    if (c <= 0) c = msm_pick_c(n);
    fe *sc = (fe *)malloc(sizeof(fe) * n);
    conv_ctx cv = {scalars, sc}; par_for(n, threads, from_mont_range, &cv);
    int nwin = (254 + c - 1) / c, nchunk = 1;
    while (nwin * nchunk < threads && (n / (nchunk + 1)) > ((uint64_t)4 << c)) nchunk++;
    g2_jac *partial = (g2_jac *)malloc(sizeof(g2_jac) * nwin * nchunk);
    msm_ctx m = {pts, sc, n, c, nwin, nchunk, partial};
    par_for((size_t)nwin * nchunk, threads, msm2_task, &m);
    g2_jac acc; g2_jac_set_inf(&acc);
    for (int w = nwin - 1; w >= 0; w--) {
        for (int i = 0; i < c; i++) g2_jac_double(&acc, &acc);
        for (int ch = 0; ch < nchunk; ch++) g2_jac_add(&acc, &acc, &partial[w * nchunk + ch]);
    }
    g2_to_aff(out, &acc); free(partial); free(sc);
}
/* out = [k] p with double-and-add (independent of the bucket method; k Montgomery Fr) */
void orc_g1_mul(const g1_aff *p, const fe *k, g1_aff *out) {
    fe kk; f_from_mont(&FR, &kk, k); g1_jac j, r; g1_from_aff(&j, p); g1_jac_mul(&r, &j, &kk); g1_to_aff(out, &r);
}
void orc_g2_mul(const g2_aff *p, const fe *k, g2_aff *out) {
    fe kk; f_from_mont(&FR, &kk, k); g2_jac j, r; g2_from_aff(&j, p); g2_jac_mul(&r, &j, &kk); g2_to_aff(out, &r);
}
void orc_g1_add(const g1_aff *a, const g1_aff *b, g1_aff *out) {
    g1_jac j, r; g1_from_aff(&j, a); g1_jac_add_mixed(&r, &j, b); g1_to_aff(out, &r);
}
void orc_g2_add(const g2_aff *a, const g2_aff *b, g2_aff *out) {
    g2_jac j, r; g2_from_aff(&j, a); g2_jac_add_mixed(&r, &j, b); g2_to_aff(out, &r);
}
int orc_g1_on_curve(const g1_aff *p) {
    if (g1_aff_is_inf(p)) return 1;
    fp_t l, r, b; fp_sqr(&l, &p->y); fp_sqr(&r, &p->x); fp_mul(&r, &r, &p->x);
    FPF(set_u64)(&FP, &b, G1_B); fp_add(&r, &r, &b); return fp_eq(&l, &r);
}

/* ------------------------------------------------------------ fft.Domain (prove.go:356-399) */
/* primitive 2^FR_ADICITY-th root of unity FR_COSET_GEN^((r-1)/2^FR_ADICITY), plain integer (gnark-crypto's fft.Domain generator) */
#if ORC_CURVE == 254
static const fe ROOT_2_ADIC = {{0x9bd61b6e725b19f0ULL, 0x402d111e41112ed4ULL, 0x00e0a7eb8ef62abcULL, 0x2a3c09f0a58a7e85ULL}};
#else
static const fe ROOT_2_ADIC = {{0x476ef4a4ec2a895eULL, 0x9b506ee363e3f04aULL, 0x60c69477d1a8a12fULL, 0x11d4b7f60cb92cc1ULL}};
#endif
static void fr_pow_u64(fe *r, const fe *a, uint64_t e) { fe ee = {{e, 0, 0, 0}}; f_pow(&FR, r, a, &ee); }
static void domain_gen(fe *w, int logn) {           /* primitive 2^logn-th root, Montgomery */
    fe g; f_to_mont(&FR, &g, &ROOT_2_ADIC);
    for (int i = logn; i < FR_ADICITY; i++) f_sqr(&FR, &g, &g);
    *w = g;
}
static size_t brev(size_t i, int logn) { size_t r = 0; for (int k = 0; k < logn; k++) { r = (r << 1) | (i & 1); i >>= 1; } return r; }

typedef struct { fe *a; const fe *tw; size_t n, half; int dit; } st_ctx;
/* one butterfly stage over butterflies [lo,hi); tw[j] = w^j, j < n/2 */
static void stage_range(size_t lo, size_t hi, int tid, void *vc) {
    (void)tid; st_ctx *s = (st_ctx *)vc;
    size_t half = s->half, step = (s->n / 2) / half;
    int lh = __builtin_ctzll(half);
    for (size_t k = lo; k < hi; k++) {
        size_t grp = k >> lh, j = k & (half - 1), i0 = (grp << (lh + 1)) + j, i1 = i0 + half;
        fe u = s->a[i0], v = s->a[i1];
        if (s->dit) {
            f_mul(&FR, &v, &v, &s->tw[j * step]);
            f_add(&FR, &s->a[i0], &u, &v); f_sub(&FR, &s->a[i1], &u, &v);
        } else {
            f_add(&FR, &s->a[i0], &u, &v); f_sub(&FR, &v, &u, &v);
            f_mul(&FR, &s->a[i1], &v, &s->tw[j * step]);
        }
    }
}
static fe *twiddles(const fe *w, size_t n) {
    size_t h = n / 2 ? n / 2 : 1;
    fe *t = (fe *)malloc(sizeof(fe) * h); t[0] = FR.one;
    for (size_t i = 1; i < h; i++) f_mul(&FR, &t[i], &t[i - 1], w);
    return t;
}
typedef struct { fe *a; fe base, scale; int logn; int bitrev_index; } sc_ctx;
/* a[pos(i)] *= scale * base^i, pos = bitrev(i) or i */
static void scale_range(size_t lo, size_t hi, int tid, void *vc) {
    (void)tid; sc_ctx *s = (sc_ctx *)vc;
    fe p; fr_pow_u64(&p, &s->base, lo); f_mul(&FR, &p, &p, &s->scale);
    for (size_t i = lo; i < hi; i++) {
        size_t pos = s->bitrev_index ? brev(i, s->logn) : i;
        f_mul(&FR, &s->a[pos], &s->a[pos], &p); f_mul(&FR, &p, &p, &s->base);
    }
}
/* mode 0: FFTInverse(DIF)  1: FFT(DIT,OnCoset)  2: FFTInverse(DIF,OnCoset)  3: FFT(DIT)  4: FFT(DIF) no scaling */
void orc_fft(fe *a, uint64_t n, int mode, int threads) {
    int logn = 0; while (((uint64_t)1 << logn) < n) logn++;
    fe w, winv, five, fiveinv, ninv, one = FR.one;
    domain_gen(&w, logn); f_inv(&FR, &winv, &w);
    f_set_u64(&FR, &five, FR_COSET_GEN); f_inv(&FR, &fiveinv, &five);
    f_set_u64(&FR, &ninv, n); f_inv(&FR, &ninv, &ninv);
    int inverse = (mode == 0 || mode == 2), dit = (mode == 1 || mode == 3);
    if (mode == 1) { sc_ctx s = {a, five, one, logn, 1}; par_for(n, threads, scale_range, &s); }
    if (n > 1) {
        fe *tw = twiddles(inverse ? &winv : &w, n);
        if (dit) for (size_t half = 1; half < n; half <<= 1) { st_ctx s = {a, tw, n, half, 1}; par_for(n / 2, threads, stage_range, &s); }
        else for (size_t half = n / 2; half >= 1; half >>= 1) { st_ctx s = {a, tw, n, half, 0}; par_for(n / 2, threads, stage_range, &s); }
        free(tw);
    }
    if (mode == 0) { sc_ctx s = {a, one, ninv, logn, 0}; par_for(n, threads, scale_range, &s); }
    if (mode == 2) { sc_ctx s = {a, fiveinv, ninv, logn, 1}; par_for(n, threads, scale_range, &s); }
}
typedef struct { fe *a; const fe *b, *c; fe den; } pw_ctx;
static void pw_range(size_t lo, size_t hi, int tid, void *vc) {
    (void)tid; pw_ctx *p = (pw_ctx *)vc;
    for (size_t i = lo; i < hi; i++) { fe t; f_mul(&FR, &t, &p->a[i], &p->b[i]); f_sub(&FR, &t, &t, &p->c[i]); f_mul(&FR, &p->a[i], &t, &p->den); }
}
/* computeH: a,b,c have nc elements (Montgomery); h receives n elements, bit-reversed order */
void orc_compute_h(const fe *a, const fe *b, const fe *c, uint64_t nc, uint64_t n, fe *h, int threads) {
    fe *bb = (fe *)calloc(n, sizeof(fe)), *cc = (fe *)calloc(n, sizeof(fe));
    memset(h, 0, n * sizeof(fe));
    memcpy(h, a, nc * sizeof(fe)); memcpy(bb, b, nc * sizeof(fe)); memcpy(cc, c, nc * sizeof(fe));
    orc_fft(h, n, 0, threads); orc_fft(bb, n, 0, threads); orc_fft(cc, n, 0, threads);
    orc_fft(h, n, 1, threads); orc_fft(bb, n, 1, threads); orc_fft(cc, n, 1, threads);
    pw_ctx p = {h, bb, cc, FR.one};
    fe five; f_set_u64(&FR, &five, FR_COSET_GEN); fr_pow_u64(&p.den, &five, n); f_sub(&FR, &p.den, &p.den, &FR.one); f_inv(&FR, &p.den, &p.den);
    par_for(n, threads, pw_range, &p);
    orc_fft(h, n, 2, threads);
    free(bb); free(cc);
}

/* ------------------------------------------------------------ MiMC-shaped chain (oracle/r1cs.py mimc_chain)
 * wires: 0=ONE 1=OUT 2=X0 3=K, then one internal wire per constraint but the last; nb_wires = nc+3.
 * A round is x <- (x + k + c_i)^e: t*t, CHAIN_SQ further squarings, one product with t.  BN254: e = 5, 3 constraints
 * (std/hash/mimc/encrypt.go:124-128,147-153); BLS12-377: e = 17, 5 constraints (encrypt.go:137-143). */
#if ORC_CURVE == 254
#define CHAIN_SQ 1
#else
#define CHAIN_SQ 3
#endif
#define CHAIN_PER_ROUND (CHAIN_SQ + 2)
typedef struct { uint64_t nc, rounds; xo_t g; uint64_t j, xw, next; } chain_it;
static void chain_begin(chain_it *it, uint64_t nc, uint64_t seed) {
    it->nc = nc; it->rounds = (nc - 1) / CHAIN_PER_ROUND; xo_seed(&it->g, seed); it->j = 0; it->xw = 2; it->next = 4;
}
/* the (nc-1)/CHAIN_PER_ROUND round constants of the chain, Montgomery (tests build the constraint matrices from them) */
void orc_mimc_chain_constants(uint64_t nc, uint64_t seed, fe *out) {
    chain_it it; chain_begin(&it, nc, seed);
    for (uint64_t r = 0; r < it.rounds; r++) xo_field(&it.g, &FR, &out[r]);
}
int orc_mimc_chain_per_round(void) { return CHAIN_PER_ROUND; }
/* W (nc+3), a,b,c (nc): solved witness and per-constraint L.w, R.w, O.w; all Montgomery */
void orc_mimc_chain_witness(uint64_t nc, uint64_t seed, const fe *x0, const fe *k, fe *W, fe *a, fe *b, fe *c) {
    chain_it it; chain_begin(&it, nc, seed);
    W[0] = FR.one; memset(&W[1], 0, sizeof(fe)); W[2] = *x0; W[3] = *k;
    uint64_t j = 0;
    for (uint64_t r = 0; r < it.rounds; r++) {
        fe ci, t; xo_field(&it.g, &FR, &ci);
        f_add(&FR, &t, &W[it.xw], k); f_add(&FR, &t, &t, &ci);
        uint64_t prev = it.next++;
        f_mul(&FR, &W[prev], &t, &t); a[j] = t; b[j] = t; c[j] = W[prev]; j++;
        for (int q = 0; q < CHAIN_SQ; q++) {
            uint64_t v = it.next++;
            f_mul(&FR, &W[v], &W[prev], &W[prev]); a[j] = W[prev]; b[j] = W[prev]; c[j] = W[v]; j++;
            prev = v;
        }
        uint64_t nx = it.next++;
        f_mul(&FR, &W[nx], &W[prev], &t); a[j] = W[prev]; b[j] = t; c[j] = W[nx]; j++;
        it.xw = nx;
    }
    while (j < nc - 1) {
        uint64_t nx = it.next++;
        f_mul(&FR, &W[nx], &W[it.xw], &W[it.xw]); a[j] = W[it.xw]; b[j] = W[it.xw]; c[j] = W[nx]; j++;
        it.xw = nx;
    }
    f_add(&FR, &W[1], &W[it.xw], k); a[j] = W[1]; b[j] = FR.one; c[j] = W[1];
}
static void acc(fe *dst, const fe *v) { f_add(&FR, dst, dst, v); }
/* trapdoor setup scalars for the chain (setup.go:143-210,364-446).
 * tw = {t, alpha, beta, gamma, delta} Montgomery.  A,B,K have nc+3 entries (K already divided by
 * gamma for public wires i<2, by delta otherwise); Z has n entries in NATURAL order. */
void orc_mimc_chain_setup_scalars(uint64_t nc, uint64_t n, uint64_t seed, const fe *tw, fe *A, fe *B, fe *K, fe *Z, int threads) {
    (void)threads;
    uint64_t nw = nc + 3; int logn = 0; while (((uint64_t)1 << logn) < n) logn++;
    fe *Cc = (fe *)calloc(nw, sizeof(fe)); memset(A, 0, nw * sizeof(fe)); memset(B, 0, nw * sizeof(fe));
    /* L_j(t) = (t^n - 1)/n * w^j / (t - w^j) */
    fe w, tn, coef, ninv; domain_gen(&w, logn);
    fr_pow_u64(&tn, &tw[0], n); f_sub(&FR, &tn, &tn, &FR.one);
    f_set_u64(&FR, &ninv, n); f_inv(&FR, &ninv, &ninv); f_mul(&FR, &coef, &tn, &ninv);
    fe *L = (fe *)malloc(sizeof(fe) * nc), *pre = (fe *)malloc(sizeof(fe) * nc), *wj = (fe *)malloc(sizeof(fe) * nc);
    fe p = FR.one, run = FR.one;
    for (uint64_t j = 0; j < nc; j++) { wj[j] = p; f_sub(&FR, &L[j], &tw[0], &p); pre[j] = run; f_mul(&FR, &run, &run, &L[j]); f_mul(&FR, &p, &p, &w); }
    fe ri; f_inv(&FR, &ri, &run);
    for (uint64_t j = nc; j-- > 0;) { fe di; f_mul(&FR, &di, &ri, &pre[j]); f_mul(&FR, &ri, &ri, &L[j]); f_mul(&FR, &di, &di, &wj[j]); f_mul(&FR, &L[j], &di, &coef); }
    free(pre); free(wj);
    chain_it it; chain_begin(&it, nc, seed);
    uint64_t j = 0;
    for (uint64_t r = 0; r < it.rounds; r++) {
        fe ci, cl; xo_field(&it.g, &FR, &ci);
        uint64_t prev = it.next++;
        f_mul(&FR, &cl, &ci, &L[j]);
        acc(&A[it.xw], &L[j]); acc(&A[3], &L[j]); acc(&A[0], &cl); acc(&B[it.xw], &L[j]); acc(&B[3], &L[j]); acc(&B[0], &cl); acc(&Cc[prev], &L[j]); j++;
        for (int q = 0; q < CHAIN_SQ; q++) {
            uint64_t v = it.next++;
            acc(&A[prev], &L[j]); acc(&B[prev], &L[j]); acc(&Cc[v], &L[j]); j++;
            prev = v;
        }
        uint64_t nx = it.next++;
        f_mul(&FR, &cl, &ci, &L[j]);
        acc(&A[prev], &L[j]); acc(&B[it.xw], &L[j]); acc(&B[3], &L[j]); acc(&B[0], &cl); acc(&Cc[nx], &L[j]); j++;
        it.xw = nx;
    }
    while (j < nc - 1) { uint64_t nx = it.next++; acc(&A[it.xw], &L[j]); acc(&B[it.xw], &L[j]); acc(&Cc[nx], &L[j]); j++; it.xw = nx; }
    acc(&A[it.xw], &L[j]); acc(&A[3], &L[j]); acc(&B[0], &L[j]); acc(&Cc[1], &L[j]);
    free(L);
    fe gi, di; f_inv(&FR, &gi, &tw[3]); f_inv(&FR, &di, &tw[4]);
    for (uint64_t i = 0; i < nw; i++) {
        fe t0, t1; f_mul(&FR, &t1, &A[i], &tw[2]); f_mul(&FR, &t0, &B[i], &tw[1]);
        f_add(&FR, &t1, &t1, &t0); f_add(&FR, &t1, &t1, &Cc[i]); f_mul(&FR, &K[i], &t1, i < 2 ? &gi : &di);
    }
    fe zdt; f_mul(&FR, &zdt, &tn, &di);
    for (uint64_t i = 0; i < n; i++) { Z[i] = zdt; f_mul(&FR, &zdt, &zdt, &tw[0]); }
    free(Cc);
}
