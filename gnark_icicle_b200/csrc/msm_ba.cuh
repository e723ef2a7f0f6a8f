// Batch-affine pre-reduction of the sorted bucket entries (included by msm.cuh).
//
// The XYZZ mixed addition of k_accumulate costs 10 base-field multiplications per entry (8M + 2S; 30 for G2).  Adding two
// AFFINE points costs 1 inversion + 2M + 1S, and inversions can be shared (Montgomery's trick: 3M per element plus one
// inversion per batch): 6M per addition (17 instead of 28 Fp multiplications for G2).  What a bucket needs is the sum of
// its entries in any order, so the entries are added in ROUNDS of independent pairs: a round turns the n entries of a
// bucket into ceil(n/2) affine points (pairs (2j, 2j+1); an odd one out is copied), all pairs of all buckets share one
// inversion tree, and after R rounds 1/2^R of the entries is left for the XYZZ kernel, which then reads a dense,
// bucket-ordered point array instead of gathering from the table.
//
// One round = k_ba_counts + scan (bucket offsets of the output), k_ba_fwd (per-thread prefix products of the pair
// denominators), the inversion tree over the per-thread totals (k_ba_up x levels, k_ba_leaf, k_ba_down x levels), and
// k_ba_bwd (unwinds the prefixes into the individual inverses and emits the sums).  Denominator of a pair: x2 - x1; 2*y1
// when the two points are equal (doubling: equal bases occur in DummySetup-like keys); 1 when a point is at infinity or
// the two cancel (the result is then the other point / infinity) -- never zero, so the shared product stays invertible.
// Over Fp2 the shared product is taken over the NORMS of the denominators (1/d = conj(d) / (a0^2 + NR a1^2)): prefixes, the
// inversion tree and the leaf inversions then work on Fp elements, and an addition costs 15 Fp multiplications instead of
// 17 (BaNorm below).
// Points in memory are canonical gnark images ((0,0) = infinity) in every round: results are bit-identical to any other
// summation order, which the existing oracle comparisons of the MSM tests check.
//
// Replaces nothing in the reference by name -- it is the inside of iciclegnark.MsmOnDevice / G1Jac.MultiExp
// (backend/groth16/bn254/icicle/icicle.go:302-382, prove.go:204-293); gnark-crypto's own CPU MultiExp uses the same
// batch-affine idea for its buckets [recollection].
#pragma once

namespace g16 {

constexpr int BA_K = 16;            // output points (pairs) per thread of k_ba_fwd / k_ba_bwd
constexpr int BA_G = 32;            // fan-in of the inversion tree
constexpr int BA_THREADS = 128;
constexpr uint32_t BA_PAIR = 0x80000000u;

// cnt_out[b] = ceil(cnt_in[b] / 2)
static __global__ void k_ba_counts(const uint32_t* __restrict__ off_in, uint32_t nb, uint32_t* __restrict__ cnt_out) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < nb) cnt_out[b] = (off_in[b + 1] - off_in[b] + 1) >> 1;
}

template <class C>
__device__ __forceinline__ typename C::E ba_load_e(const typename C::E* p) {
    typename C::E r;
    constexpr int Q = sizeof(typename C::E) / 16;
    const uint4* s = reinterpret_cast<const uint4*>(p);
    uint4* d = reinterpret_cast<uint4*>(&r);
#pragma unroll
    for (int i = 0; i < Q; i++) d[i] = s[i];
    return r;
}
// the same through the read-only path with a 64-byte L2 fill: for the random gathers of round 0 (ncu, 2^22-point G1 MSM:
// 7.06 GB read by k_ba_fwd for 54.5 M 32-byte coordinates -- every access filled a 128-byte line)
template <class C>
__device__ __forceinline__ typename C::E ba_gather_e(const typename C::E* p) {
    typename C::E r;
    constexpr int Q = sizeof(typename C::E) / 16;
    const uint4* s = reinterpret_cast<const uint4*>(p);
    uint4* d = reinterpret_cast<uint4*>(&r);
#pragma unroll
    for (int i = 0; i < Q; i++)
        asm volatile("ld.global.nc.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(d[i].x), "=r"(d[i].y), "=r"(d[i].z), "=r"(d[i].w) : "l"(s + i));
    return r;
}
template <class C>
__device__ __forceinline__ void ba_store_e(typename C::E* p, const typename C::E& v) {
    constexpr int Q = sizeof(typename C::E) / 16;
    const uint4* s = reinterpret_cast<const uint4*>(&v);
    uint4* d = reinterpret_cast<uint4*>(p);
#pragma unroll
    for (int i = 0; i < Q; i++) d[i] = s[i];
}

// the same for any 16-byte-multiple value type (elements of the norm field)
template <class T>
__device__ __forceinline__ T ba_ld(const T* p) {
    T r;
    const uint4* s = reinterpret_cast<const uint4*>(p);
    uint4* d = reinterpret_cast<uint4*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); i++) d[i] = s[i];
    return r;
}
template <class T>
__device__ __forceinline__ void ba_st(T* p, const T& v) {
    const uint4* s = reinterpret_cast<const uint4*>(&v);
    uint4* d = reinterpret_cast<uint4*>(p);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); i++) d[i] = s[i];
}

// The field the shared inversion runs in.  Prime base field: the denominators themselves.  Fp2: their norms
// a0^2 + NR a1^2 in Fp (2 squarings in the forward pass, which also stores them for the backward pass); the inverse of
// the denominator is conj(d) * norm^-1 (2 multiplications) -- 1 + 1 + 2 = 4 Fp multiplications in the backward pass where
// the Fp2 products took 6, and a third of the work in the tree.
template <class BF>
struct BaNorm {
    typedef BF NF;
    static constexpr bool EXT = false;
    __device__ __forceinline__ static typename BF::T norm(const typename BF::T& d) { return d; }
    __device__ __forceinline__ static typename BF::T inverse(const typename BF::T&, const typename BF::T& ninv) { return ninv; }
};
template <class F, int NR>
struct BaNorm<Ext2<F, NR>> {
    typedef F NF;
    typedef typename Ext2<F, NR>::T E;
    static constexpr bool EXT = true;
    __device__ __forceinline__ static typename F::T norm(const E& d) { return F::add(F::sqr(d.a0), Ext2<F, NR>::mul_nr(F::sqr(d.a1))); }
    __device__ __forceinline__ static E inverse(const E& d, const typename F::T& ninv) {
        E r; r.a0 = F::mul(d.a0, ninv); r.a1 = F::neg(F::mul(d.a1, ninv)); return r;
    }
};
template <class F, int NR>
struct BaNorm<Ext2Call<F, NR>> : BaNorm<Ext2<F, NR>> {};

// where the inputs of a round come from: round 0 gathers table points through the sorted entries (sign in bit 31),
// later rounds read the previous round's dense output
template <class C>
struct BaSrc {
    const typename C::Aff* pts;
    const uint32_t* sorted;     // nullptr: direct
    __device__ __forceinline__ const typename C::Aff* addr(uint32_t i, uint32_t& neg) const {
        if (sorted) { const uint32_t e = sorted[i]; neg = e >> 31; return pts + (e & 0x7fffffffu); }
        neg = 0;
        return pts + i;
    }
};

// classification of a pair; d is the denominator that goes into the shared product
enum { BA_FIRST_INF = 0, BA_SECOND_INF = 1, BA_CANCEL = 2, BA_DOUBLE = 3, BA_ADD = 4 };

template <class C>
__device__ __forceinline__ int ba_classify(const typename C::Aff& p1, const typename C::Aff& p2, typename C::E& d) {
    typedef typename C::BaseField BF;
    if (C::is_inf(p1)) return BA_FIRST_INF;
    if (C::is_inf(p2)) return BA_SECOND_INF;
    d = BF::sub(p2.x, p1.x);
    if (BF::is_zero(d)) {
        if (!BF::eq(p1.y, p2.y)) return BA_CANCEL;            // same x, different y: P + (-P)
        d = BF::dbl(p1.y);
        if (BF::is_zero(d)) return BA_CANCEL;                  // a point of order two doubles to infinity
        return BA_DOUBLE;
    }
    return BA_ADD;
}

template <class C>
__device__ __forceinline__ typename C::Aff ba_load_pt(const BaSrc<C>& src, uint32_t i) {
    uint32_t neg;
    const typename C::Aff* a = src.addr(i, neg);
    typename C::Aff p = load_aff<C>(a);
    if (neg) p = C::neg(p);
    return p;
}
// the forward pass needs the x coordinates only (half the bytes of the gather); y is fetched in the rare cases that
// need it: x = 0 (a point at infinity?) or equal x (doubling or cancellation?)
template <class C>
__device__ __forceinline__ int ba_classify_x(const BaSrc<C>& src, uint32_t i0, typename C::E& d) {
    typedef typename C::BaseField BF;
    typedef typename C::E E;
    uint32_t n1, n2;
    const typename C::Aff* a1 = src.addr(i0, n1);
    const typename C::Aff* a2 = src.addr(i0 + 1, n2);
    const E x1 = src.sorted ? ba_gather_e<C>(&a1->x) : ba_load_e<C>(&a1->x), x2 = src.sorted ? ba_gather_e<C>(&a2->x) : ba_load_e<C>(&a2->x);
    if (BF::is_zero(x1) || BF::is_zero(x2) || BF::eq(x1, x2)) {
        typename C::Aff p1, p2;
        p1.x = x1; p1.y = ba_load_e<C>(&a1->y); if (n1) p1 = C::neg(p1);
        p2.x = x2; p2.y = ba_load_e<C>(&a2->y); if (n2) p2 = C::neg(p2);
        return ba_classify<C>(p1, p2, d);
    }
    d = BF::sub(x2, x1);
    return BA_ADD;
}

// Work assignment: a warp owns 32 K consecutive output points, lane l the points base + 32 j + l (j < K), so that every
// access to the per-output arrays (descriptor, prefix, result) is coalesced across the warp.
// k_ba_fwd: descriptor (first input index | pair flag) of every output, running product of the lane's pair denominators
// after every output, and the product of all of them per thread.
template <class C>
__global__ void __launch_bounds__(BA_THREADS)
k_ba_fwd(BaSrc<C> src, const uint32_t* __restrict__ off_in, const uint32_t* __restrict__ off_out, uint32_t nb,
         uint32_t* __restrict__ desc, typename BaNorm<typename C::BaseField>::NF::T* __restrict__ prefix,
         typename BaNorm<typename C::BaseField>::NF::T* __restrict__ nrm, typename BaNorm<typename C::BaseField>::NF::T* __restrict__ tot,
         uint32_t nthreads) {
    typedef BaNorm<typename C::BaseField> BN;
    typedef typename BN::NF NF;
    typedef typename NF::T NE;
    typedef typename C::E E;
    const uint32_t t = blockIdx.x * BA_THREADS + threadIdx.x;
    if (t >= nthreads) return;
    const uint32_t total = off_out[nb];
    const uint64_t o0 = (uint64_t)(t >> 5) * (32 * BA_K) + (t & 31);
    NE run = NF::one();
    if (o0 < total) {
        uint32_t lo = 0, hi = nb;                         // off_out[lo] <= o0 < off_out[hi]
        while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (off_out[mid] <= (uint32_t)o0) lo = mid; else hi = mid; }
        uint32_t b = lo, b_out = off_out[b], b_next = off_out[b + 1], b_in = off_in[b], b_in_end = off_in[b + 1];
        for (int j = 0; j < BA_K; j++) {
            const uint64_t o64 = o0 + 32ull * j;
            if (o64 >= total) break;
            const uint32_t o = (uint32_t)o64;
            while (o >= b_next) { b++; b_out = b_next; b_next = off_out[b + 1]; b_in = b_in_end; b_in_end = off_in[b + 1]; }
            const uint32_t i0 = b_in + 2 * (o - b_out);
            const bool pair = i0 + 1 < b_in_end;
            desc[o] = i0 | (pair ? BA_PAIR : 0u);
            if (pair) {
                E d;
                if (ba_classify_x<C>(src, i0, d) >= BA_DOUBLE) {
                    const NE nd = BN::norm(d);
                    if (BN::EXT) ba_st<NE>(nrm + o, nd);
                    run = NF::mul(run, nd);
                }
            }
            ba_st<NE>(prefix + o, run);
        }
    }
    ba_st<NE>(tot + t, run);
}

template <class C>
__global__ void __launch_bounds__(BA_THREADS)
k_ba_bwd(BaSrc<C> src, const uint32_t* __restrict__ off_out, uint32_t nb, const uint32_t* __restrict__ desc,
         const typename BaNorm<typename C::BaseField>::NF::T* __restrict__ prefix, const typename BaNorm<typename C::BaseField>::NF::T* __restrict__ nrm,
         const typename BaNorm<typename C::BaseField>::NF::T* __restrict__ inv_tot, typename C::Aff* __restrict__ out, uint32_t nthreads) {
    typedef typename C::BaseField BF;
    typedef BaNorm<BF> BN;
    typedef typename BN::NF NF;
    typedef typename NF::T NE;
    typedef typename C::E E;
    typedef typename C::Aff Aff;
    const uint32_t t = blockIdx.x * BA_THREADS + threadIdx.x;
    if (t >= nthreads) return;
    const uint32_t total = off_out[nb];
    const uint64_t o0 = (uint64_t)(t >> 5) * (32 * BA_K) + (t & 31);
    if (o0 >= total) return;
    int jmax = BA_K - 1;
    if (o0 + 32ull * jmax >= total) jmax = (int)((total - 1 - o0) / 32);
    NE acc = ba_ld<NE>(inv_tot + t);                      // 1 / (product of this thread's denominators, or of their norms)
    for (int j = jmax; j >= 0; j--) {
        const uint32_t o = (uint32_t)(o0 + 32ull * j);
        const uint32_t de = desc[o], i0 = de & ~BA_PAIR;
        const Aff p1 = ba_load_pt<C>(src, i0);
        Aff r;
        if (!(de & BA_PAIR)) {
            r = p1;
        } else {
            const Aff p2 = ba_load_pt<C>(src, i0 + 1);
            E d;
            const int kind = ba_classify<C>(p1, p2, d);
            if (kind == BA_FIRST_INF) r = p2;
            else if (kind == BA_SECOND_INF) r = p1;
            else if (kind == BA_CANCEL) r = C::aff_inf();
            else {
                const NE ninv = j == 0 ? acc : NF::mul(acc, ba_ld<NE>(prefix + o - 32));
                if (BN::EXT) acc = NF::mul(acc, ba_ld<NE>(nrm + o)); else acc = NF::mul(acc, BN::norm(d));
                const E dinv = BN::inverse(d, ninv);
                E lam;
                if (kind == BA_DOUBLE) { const E x2 = BF::sqr(p1.x); lam = BF::mul(BF::add(BF::dbl(x2), x2), dinv); }
                else lam = BF::mul(BF::sub(p2.y, p1.y), dinv);
                r.x = BF::sub(BF::sub(BF::sqr(lam), p1.x), p2.x);
                r.y = BF::sub(BF::mul(lam, BF::sub(p1.x, r.x)), p1.y);
            }
        }
        r.x = BF::canon(r.x); r.y = BF::canon(r.y);       // memory only ever holds canonical coordinates
        uint4* dst = reinterpret_cast<uint4*>(out + o);
        const uint4* s = reinterpret_cast<const uint4*>(&r);
#pragma unroll
        for (int q = 0; q < (int)(sizeof(Aff) / 16); q++) dst[q] = s[q];
    }
}

// ---- inversion tree over an array of non-zero elements of the field NF (the norm field of the round)
// up: group u = elements [u G, (u+1) G): pref[e] = product of the group's elements up to e, out[u] = product of the group
template <class NF>
__global__ void __launch_bounds__(BA_THREADS)
k_ba_up(const typename NF::T* __restrict__ in, uint32_t n, typename NF::T* __restrict__ pref, typename NF::T* __restrict__ out, uint32_t groups) {
    typedef typename NF::T NE;
    const uint32_t u = blockIdx.x * BA_THREADS + threadIdx.x;
    if (u >= groups) return;
    NE run = NF::one();
    const uint64_t e0 = (uint64_t)u * BA_G;
    for (int k = 0; k < BA_G && e0 + k < n; k++) {
        run = NF::mul(run, ba_ld<NE>(in + e0 + k));
        ba_st<NE>(pref + e0 + k, run);
    }
    ba_st<NE>(out + u, run);
}
template <class NF>
__global__ void __launch_bounds__(BA_THREADS) k_ba_leaf(typename NF::T* __restrict__ a, uint32_t n) {
    typedef typename NF::T NE;
    const uint32_t i = blockIdx.x * BA_THREADS + threadIdx.x;
    // binary Euclid, not the Fermat power: this one thread's chain is the latency of the round (0.17 ms as a power)
    if (i < n) ba_st<NE>(a + i, NF::inv_binary(NF::canon(ba_ld<NE>(a + i))));
}
// down: inv_grp[u] = inverse of the group's product  ->  inv[e] for every element of the group (in and inv may alias)
template <class NF>
__global__ void __launch_bounds__(BA_THREADS)
k_ba_down(const typename NF::T* in, uint32_t n, const typename NF::T* __restrict__ pref, const typename NF::T* __restrict__ inv_grp,
          typename NF::T* inv, uint32_t groups) {
    typedef typename NF::T NE;
    const uint32_t u = blockIdx.x * BA_THREADS + threadIdx.x;
    if (u >= groups) return;
    NE acc = ba_ld<NE>(inv_grp + u);
    const uint64_t e0 = (uint64_t)u * BA_G;
    int last = BA_G - 1;
    if (e0 + last >= n) last = (int)(n - 1 - e0);
    for (int k = last; k >= 0; k--) {
        const NE x = ba_ld<NE>(in + e0 + k);
        const NE v = k ? NF::mul(acc, ba_ld<NE>(pref + e0 + k - 1)) : acc;
        ba_st<NE>(inv + e0 + k, v);
        acc = NF::mul(acc, x);
    }
}

}  // namespace g16
