// Montgomery prime fields on 32-bit limbs for sm_100a (device: mad.lo.cc / madc.hi.cc carry chains that
// ptxas fuses into IMAD.WIDE.U32.X; host: portable 64-bit code used only by the O(1) epilogue).
//
// Replaces the arithmetic the reference obtains from gnark-crypto fr/fp (go.mod:11; call sites
// backend/groth16/bn254/prove.go:356-399 for fr, :202-304 for fp) and, on its GPU path, from icicle's
// device field (backend/groth16/bn254/icicle/icicle.go:489-510 NTT/vec ops, :302-382 MSM).
// Multiplier lineage: mul_n / cmad_n / madc_n_rshift / mad_redc below are the even/odd-accumulator CIOS layout for
// 32-bit limbs published in supranational/sppark (ff/mont_t.cuh) and reused by ICICLE's device field header: two
// interleaved accumulators so that every mad.lo.cc/madc.hi.cc pair continues one carry chain.  Neither project is
// in /root/reference; the layout is restated here from the published technique, the code is this repository's.
// Memory image = gnark-crypto's: N/2 little-endian uint64 limbs == N little-endian uint32 limbs,
// Montgomery form with R = 2^(32N).
#pragma once
#include "curve_select.h"
#include <cstdint>
#include <cstring>

#ifdef __CUDACC__
#define G16_HD __host__ __device__ __forceinline__
#define G16_D __device__ __forceinline__
// cold paths (inversion, full group additions, doublings): real calls keep compile time and
// register pressure of the hot kernels down
#define G16_HD_COLD __host__ __device__ __noinline__
#else
#define G16_HD inline
#define G16_D inline
#define G16_HD_COLD inline
#endif

namespace g16 {

template <int N>
struct alignas(16) Big {
    uint32_t l[N];
};

// ---------------------------------------------------------------- field parameter packs
// mod(i) etc. are switch-based so that, after unrolling, every limb is an immediate operand.
struct FpBN254 {
    static constexpr int N = 8;
    static constexpr uint32_t INV = 0xe4866389u;  // -p^-1 mod 2^32
    G16_HD static constexpr uint32_t mod(int i) {
        constexpr uint32_t m[8] = {0xd87cfd47u, 0x3c208c16u, 0x6871ca8du, 0x97816a91u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return m[i];
    }
    G16_HD static constexpr uint32_t one(int i) {  // R mod p
        constexpr uint32_t m[8] = {0xc58f0d9du, 0xd35d438du, 0xf5c70b3du, 0x0a78eb28u, 0x7879462cu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return m[i];
    }
    G16_HD static constexpr uint32_t r2(int i) {  // R^2 mod p
        constexpr uint32_t m[8] = {0x538afa89u, 0xf32cfc5bu, 0xd44501fbu, 0xb5e71911u, 0x0a417ff6u, 0x47ab1effu, 0xcab8351fu, 0x06d89f71u};
        return m[i];
    }
};
struct FrBN254 {
    static constexpr int N = 8;
    static constexpr uint32_t INV = 0xefffffffu;
    G16_HD static constexpr uint32_t mod(int i) {
        constexpr uint32_t m[8] = {0xf0000001u, 0x43e1f593u, 0x79b97091u, 0x2833e848u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return m[i];
    }
    G16_HD static constexpr uint32_t one(int i) {
        constexpr uint32_t m[8] = {0x4ffffffbu, 0xac96341cu, 0x9f60cd29u, 0x36fc7695u, 0x7879462eu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return m[i];
    }
    G16_HD static constexpr uint32_t r2(int i) {
        constexpr uint32_t m[8] = {0xae216da7u, 0x1bb8e645u, 0xe35c59e3u, 0x53fe3ab1u, 0x53bb8085u, 0x8c49833du, 0x7f4e44a5u, 0x0216d0b1u};
        return m[i];
    }
};

// BLS12-377 (BASELINE config 5): 377-bit Fp on 12 limbs, 253-bit Fr on 8 limbs; both moduli are 1 mod 2^32.
struct FpBLS377 {
    static constexpr int N = 12;
    static constexpr uint32_t INV = 0xffffffffu;
    G16_HD static constexpr uint32_t mod(int i) {
        constexpr uint32_t m[12] = {0x00000001u, 0x8508c000u, 0x30000000u, 0x170b5d44u, 0xba094800u, 0x1ef3622fu, 0x00f5138fu, 0x1a22d9f3u, 0x6ca1493bu, 0xc63b05c0u, 0x17c510eau, 0x01ae3a46u};
        return m[i];
    }
    G16_HD static constexpr uint32_t one(int i) {
        constexpr uint32_t m[12] = {0xffffff68u, 0x02cdffffu, 0x7fffffb1u, 0x51409f83u, 0x8a7d3ff2u, 0x9f7db3a9u, 0x6e7c6305u, 0x7b4e97b7u, 0x803c84e8u, 0x4cf495bfu, 0xe2fdf49au, 0x008d6661u};
        return m[i];
    }
    G16_HD static constexpr uint32_t r2(int i) {
        constexpr uint32_t m[12] = {0x9400cd22u, 0xb786686cu, 0xb00431b1u, 0x0329fcaau, 0x62d6b46du, 0x22a5f111u, 0x827dc3acu, 0xbfdf7d03u, 0x41790bf9u, 0x837e92f0u, 0x1e914b88u, 0x006dfccbu};
        return m[i];
    }
};
struct FrBLS377 {
    static constexpr int N = 8;
    static constexpr uint32_t INV = 0xffffffffu;
    G16_HD static constexpr uint32_t mod(int i) {
        constexpr uint32_t m[8] = {0x00000001u, 0x0a118000u, 0xd0000001u, 0x59aa76feu, 0x5c37b001u, 0x60b44d1eu, 0x9a2ca556u, 0x12ab655eu};
        return m[i];
    }
    G16_HD static constexpr uint32_t one(int i) {
        constexpr uint32_t m[8] = {0xfffffff3u, 0x7d1c7fffu, 0x6ffffff2u, 0x7257f50fu, 0x512c0feeu, 0x16d81575u, 0x2bbb9a9du, 0x0d4bda32u};
        return m[i];
    }
    G16_HD static constexpr uint32_t r2(int i) {
        constexpr uint32_t m[8] = {0xb861857bu, 0x25d577bau, 0x8860591fu, 0xcc2c27b5u, 0xe5dc8593u, 0xa7cc008fu, 0xeff1c939u, 0x011fdae7u};
        return m[i];
    }
};

#ifdef __CUDA_ARCH__
// ---------------------------------------------------------------- PTX carry-chain primitives
namespace ptx {
G16_D uint32_t add_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
G16_D uint32_t addc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
G16_D uint32_t addc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
G16_D uint32_t sub_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
G16_D uint32_t subc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
G16_D uint32_t subc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
G16_D uint32_t mul_lo(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
G16_D uint32_t mul_hi(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
G16_D uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
G16_D uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
G16_D uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
G16_D uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
}  // namespace ptx
#endif

// ---------------------------------------------------------------- the field
template <class P>
struct Field {
    static constexpr int N = P::N;
    typedef Big<N> T;

    G16_HD static T zero() { T r;
#pragma unroll
        for (int i = 0; i < N; i++) r.l[i] = 0; return r; }
    G16_HD static T one() { T r;
#pragma unroll
        for (int i = 0; i < N; i++) r.l[i] = P::one(i); return r; }
    G16_HD static T r2() { T r;
#pragma unroll
        for (int i = 0; i < N; i++) r.l[i] = P::r2(i); return r; }
    G16_HD static T modulus() { T r;
#pragma unroll
        for (int i = 0; i < N; i++) r.l[i] = P::mod(i); return r; }
    G16_HD static bool is_zero(const T& a) { uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < N; i++) o |= a.l[i]; return o == 0; }
    G16_HD static bool eq(const T& a, const T& b) { uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < N; i++) o |= a.l[i] ^ b.l[i]; return o == 0; }

    G16_HD static T canon(const T& a) { return a; }   // values of this field are always in [0, p); see LazyFieldCall
    // r = a - p if a >= p (a < 2p)
    G16_HD static T reduce_once(const T& a) {
        T s;
#ifdef __CUDA_ARCH__
        s.l[0] = ptx::sub_cc(a.l[0], P::mod(0));
#pragma unroll
        for (int i = 1; i < N; i++) s.l[i] = ptx::subc_cc(a.l[i], P::mod(i));
        uint32_t borrow = ptx::subc(0, 0);  // 0xffffffff if a < p
        T r;
#pragma unroll
        for (int i = 0; i < N; i++) r.l[i] = borrow ? a.l[i] : s.l[i];
        return r;
#else
        uint64_t br = 0;
        for (int i = 0; i < N; i++) { uint64_t d = (uint64_t)a.l[i] - P::mod(i) - br; s.l[i] = (uint32_t)d; br = (d >> 32) & 1; }
        return br ? a : s;
#endif
    }

    G16_HD static T add(const T& a, const T& b) {
        T r;
#ifdef __CUDA_ARCH__
        r.l[0] = ptx::add_cc(a.l[0], b.l[0]);
#pragma unroll
        for (int i = 1; i < N - 1; i++) r.l[i] = ptx::addc_cc(a.l[i], b.l[i]);
        r.l[N - 1] = ptx::addc(a.l[N - 1], b.l[N - 1]);  // p < 2^(32N-1): no carry out
#else
        uint64_t c = 0;
        for (int i = 0; i < N; i++) { c += (uint64_t)a.l[i] + b.l[i]; r.l[i] = (uint32_t)c; c >>= 32; }
#endif
        return reduce_once(r);
    }
    G16_HD static T dbl(const T& a) { return add(a, a); }

    G16_HD static T sub(const T& a, const T& b) {
        T r;
#ifdef __CUDA_ARCH__
        r.l[0] = ptx::sub_cc(a.l[0], b.l[0]);
#pragma unroll
        for (int i = 1; i < N; i++) r.l[i] = ptx::subc_cc(a.l[i], b.l[i]);
        uint32_t borrow = ptx::subc(0, 0);
        r.l[0] = ptx::add_cc(r.l[0], borrow & P::mod(0));
#pragma unroll
        for (int i = 1; i < N - 1; i++) r.l[i] = ptx::addc_cc(r.l[i], borrow & P::mod(i));
        r.l[N - 1] = ptx::addc(r.l[N - 1], borrow & P::mod(N - 1));
#else
        uint64_t br = 0;
        for (int i = 0; i < N; i++) { uint64_t d = (uint64_t)a.l[i] - b.l[i] - br; r.l[i] = (uint32_t)d; br = (d >> 32) & 1; }
        if (br) { uint64_t c = 0; for (int i = 0; i < N; i++) { c += (uint64_t)r.l[i] + P::mod(i); r.l[i] = (uint32_t)c; c >>= 32; } }
#endif
        return r;
    }
    G16_HD static T neg(const T& a) { return is_zero(a) ? a : sub(zero(), a); }

#ifdef __CUDA_ARCH__
    // acc[j], acc[j+1] = a[j] * bi for even j  (a may point at the odd limbs)
    G16_D static void mul_n(uint32_t* acc, const uint32_t* a, uint32_t bi) {
#pragma unroll
        for (int j = 0; j < N; j += 2) { acc[j] = ptx::mul_lo(a[j], bi); acc[j + 1] = ptx::mul_hi(a[j], bi); }
    }
    // acc += (a[0], a[2], ...) * bi as one carry chain; carry-out is left in CC
    G16_D static void cmad_n(uint32_t* acc, const uint32_t* a, uint32_t bi) {
        acc[0] = ptx::mad_lo_cc(a[0], bi, acc[0]);
        acc[1] = ptx::madc_hi_cc(a[0], bi, acc[1]);
#pragma unroll
        for (int j = 2; j < N; j += 2) { acc[j] = ptx::madc_lo_cc(a[j], bi, acc[j]); acc[j + 1] = ptx::madc_hi_cc(a[j], bi, acc[j + 1]); }
    }
    // same with the modulus limbs (immediates): acc += (p[o], p[o+2], ...) * m
    template <int O>
    G16_D static void cmad_mod(uint32_t* acc, uint32_t m) {
        acc[0] = ptx::mad_lo_cc(P::mod(O), m, acc[0]);
        acc[1] = ptx::madc_hi_cc(P::mod(O), m, acc[1]);
#pragma unroll
        for (int j = 2; j < N; j += 2) { acc[j] = ptx::madc_lo_cc(P::mod(O + j), m, acc[j]); acc[j + 1] = ptx::madc_hi_cc(P::mod(O + j), m, acc[j + 1]); }
    }
    // y <- (y >> 64) + (a[0], a[2], ...) * bi, consuming the carry in CC (a points at the odd limbs)
    G16_D static void madc_n_rshift(uint32_t* y, const uint32_t* a, uint32_t bi) {
#pragma unroll
        for (int j = 0; j < N - 2; j += 2) { y[j] = ptx::madc_lo_cc(a[j], bi, y[j + 2]); y[j + 1] = ptx::madc_hi_cc(a[j], bi, y[j + 3]); }
        y[N - 2] = ptx::madc_lo_cc(a[N - 2], bi, 0);
        y[N - 1] = ptx::madc_hi(a[N - 1 - 1], bi, 0);
    }
    // one interleaved multiply-reduce step; x plays "even" (position 0), y plays "odd" (position 1)
    G16_D static void mad_redc(uint32_t* x, uint32_t* y, const uint32_t* a, uint32_t bi, bool first) {
        if (first) {
            mul_n(y, a + 1, bi);
            mul_n(x, a, bi);
        } else {
            x[0] = ptx::add_cc(x[0], y[1]);
            madc_n_rshift(y, a + 1, bi);
            cmad_n(x, a, bi);
            y[N - 1] = ptx::addc(y[N - 1], 0);
        }
        uint32_t m = x[0] * P::INV;
        cmad_mod<1>(y, m);
        cmad_mod<0>(x, m);
        y[N - 1] = ptx::addc(y[N - 1], 0);
    }
#endif

#ifdef __CUDA_ARCH__
    // Montgomery product without the final conditional subtraction: for inputs in [0, 2p) the result is in [0, 2p)
    // because 4p < R for every modulus here (a*b/R < p, plus at most p from the reduction)
    G16_D static T mul_raw(const T& a, const T& b) {
        uint32_t even[N], odd[N];
#pragma unroll
        for (int i = 0; i < N; i += 2) {
            mad_redc(even, odd, a.l, b.l[i], i == 0);
            mad_redc(odd, even, a.l, b.l[i + 1], false);
        }
        T r;
        r.l[0] = ptx::add_cc(even[0], odd[1]);
#pragma unroll
        for (int i = 1; i < N - 1; i++) r.l[i] = ptx::addc_cc(even[i], odd[i + 1]);
        r.l[N - 1] = ptx::addc(even[N - 1], 0);
        return r;
    }
#endif

    // Montgomery product a*b*R^-1 mod p, inputs and output in [0, p)
    G16_HD static T mul(const T& a, const T& b) {
#ifdef __CUDA_ARCH__
        uint32_t even[N], odd[N];
#pragma unroll
        for (int i = 0; i < N; i += 2) {
            mad_redc(even, odd, a.l, b.l[i], i == 0);
            mad_redc(odd, even, a.l, b.l[i + 1], false);
        }
        // result = even + (odd >> 32)
        T r;
        r.l[0] = ptx::add_cc(even[0], odd[1]);
#pragma unroll
        for (int i = 1; i < N - 1; i++) r.l[i] = ptx::addc_cc(even[i], odd[i + 1]);
        r.l[N - 1] = ptx::addc(even[N - 1], 0);
        return reduce_once(r);
#else
        // host: CIOS on 64-bit limbs (epilogue only; never on the accelerated path)
        constexpr int M = N / 2;
        typedef unsigned __int128 u128;
        uint64_t A[M], B[M], Q[M], t[M + 2];
        for (int i = 0; i < M; i++) {
            A[i] = (uint64_t)a.l[2 * i] | ((uint64_t)a.l[2 * i + 1] << 32);
            B[i] = (uint64_t)b.l[2 * i] | ((uint64_t)b.l[2 * i + 1] << 32);
            Q[i] = (uint64_t)P::mod(2 * i) | ((uint64_t)P::mod(2 * i + 1) << 32);
        }
        uint64_t x = (uint64_t)(0u - P::INV);          // p^-1 mod 2^32
        x = x * (2 - Q[0] * x);                        // Newton step -> mod 2^64
        const uint64_t inv64 = 0 - x;
        for (int i = 0; i < M + 2; i++) t[i] = 0;
        for (int i = 0; i < M; i++) {
            u128 c = 0;
            for (int j = 0; j < M; j++) { c += (u128)A[j] * B[i] + t[j]; t[j] = (uint64_t)c; c >>= 64; }
            c += t[M]; t[M] = (uint64_t)c; t[M + 1] = (uint64_t)(c >> 64);
            uint64_t m = t[0] * inv64;
            c = (u128)m * Q[0] + t[0]; c >>= 64;
            for (int j = 1; j < M; j++) { c += (u128)m * Q[j] + t[j]; t[j - 1] = (uint64_t)c; c >>= 64; }
            c += t[M]; t[M - 1] = (uint64_t)c; t[M] = t[M + 1] + (uint64_t)(c >> 64);
        }
        T r;
        for (int i = 0; i < M; i++) { r.l[2 * i] = (uint32_t)t[i]; r.l[2 * i + 1] = (uint32_t)(t[i] >> 32); }
        return reduce_once(r);
#endif
    }
    G16_HD static T sqr(const T& a) { return mul(a, a); }
    G16_HD static T to_mont(const T& a) { return mul(a, r2()); }
    G16_HD static T from_mont(const T& a) { T o = zero(); o.l[0] = 1; return mul(a, o); }

    // a^e, e a plain little-endian integer of `words` 32-bit words (not constant time; host/tooling use)
    G16_HD_COLD static T pow(const T& a, const uint32_t* e, int words) {
        T acc = one();
        for (int i = words * 32 - 1; i >= 0; i--) {
            acc = sqr(acc);
            if ((e[i >> 5] >> (i & 31)) & 1) acc = mul(acc, a);
        }
        return acc;
    }
    G16_HD static T pow_u64(const T& a, uint64_t e) {
        uint32_t w[2] = {(uint32_t)e, (uint32_t)(e >> 32)};
        return pow(a, w, 2);
    }
    // Fermat inversion a^(p-2); inv(0) = 0
    G16_HD_COLD static T inv(const T& a) {
        uint32_t e[N];
        uint64_t br = 2;
        for (int i = 0; i < N; i++) { uint64_t d = (uint64_t)P::mod(i) - br; e[i] = (uint32_t)d; br = (d >> 32) & 1; }
        return pow(a, e, N);
    }
    G16_HD static T from_u64(uint64_t v) { T t = zero(); t.l[0] = (uint32_t)v; t.l[1] = (uint32_t)(v >> 32); return to_mont(t); }

    // The same inverse by the binary extended Euclidean algorithm (right shifts only, the modulus is odd): for the
    // inversion-tree leaves of the batch-affine rounds (msm_ba.cuh), where ONE thread's chain is the latency of the whole
    // round -- ~0.7 log2(p) iterations of limb-wise shifts, masks and carry chains instead of the ~1.5 log2(p) dependent
    // multiplications of the Fermat power.  Branch-free inside an iteration (the lanes of a warp only differ in their
    // iteration count).  Input and output canonical Montgomery images; inv_binary(0) = 0.
    // Invariants: x1 * a = u and x2 * a = v (mod p) on the plain integer a behind the image; v stays odd; u reaches 0
    // with v = gcd = 1, so x2 = a^-1.  One iteration: if u is odd, make it even (u -= v after ordering them, x1 -= x2),
    // then strip all k <= 31 trailing zero bits of u at once and divide x1 by 2^k mod p: x1 = (x1 + m p) >> k with
    // m = x1 * (-p^-1) mod 2^k (one row of a Montgomery reduction).
    // The integer handed in is aR, so x2 = a^-1 R^-1: two to_mont bring it to a^-1 R.
    G16_HD_COLD static T inv_binary(const T& a) {
        uint32_t u[N], v[N], x1[N], x2[N];
#pragma unroll
        for (int i = 0; i < N; i++) { u[i] = a.l[i]; v[i] = P::mod(i); x1[i] = 0; x2[i] = 0; }
        x1[0] = 1;
        for (int it = 0; it < 64 * N; it++) {                 // bitlen(u) + bitlen(v) drops every iteration: < 64 N of them
            uint32_t nz = 0;
#pragma unroll
            for (int i = 0; i < N; i++) nz |= u[i];
            if (!nz) break;
            const uint32_t odd = 0u - (u[0] & 1u);
            // lt = odd and u < v (borrow of u - v)
            uint32_t br = 0;
#pragma unroll
            for (int i = 0; i < N; i++) { const uint64_t d = (uint64_t)u[i] - v[i] - br; br = (uint32_t)(d >> 32) & 1u; }
            const uint32_t lt = odd & (0u - br);
#pragma unroll
            for (int i = 0; i < N; i++) {                     // conditional swap (u, v), (x1, x2)
                uint32_t t = (u[i] ^ v[i]) & lt; u[i] ^= t; v[i] ^= t;
                t = (x1[i] ^ x2[i]) & lt; x1[i] ^= t; x2[i] ^= t;
            }
            // odd: u -= v (now u >= v), x1 -= x2 mod p
            br = 0;
#pragma unroll
            for (int i = 0; i < N; i++) { const uint64_t d = (uint64_t)u[i] - (v[i] & odd) - br; u[i] = (uint32_t)d; br = (uint32_t)(d >> 32) & 1u; }
            br = 0;
#pragma unroll
            for (int i = 0; i < N; i++) { const uint64_t d = (uint64_t)x1[i] - (x2[i] & odd) - br; x1[i] = (uint32_t)d; br = (uint32_t)(d >> 32) & 1u; }
            const uint32_t fix = 0u - br;
            uint32_t cy = 0;
#pragma unroll
            for (int i = 0; i < N; i++) { const uint64_t d = (uint64_t)x1[i] + (P::mod(i) & fix) + cy; x1[i] = (uint32_t)d; cy = (uint32_t)(d >> 32); }
            // u is even now (or zero): strip k trailing zero bits, 1 <= k <= 31 (k = 0 only when u became 0: nothing moves)
            uint32_t low = u[0] | 0x80000000u;                // a zero low limb moves 31 bits now and the rest next time
            int k = 0;
#ifdef __CUDA_ARCH__
            k = __ffs((int)low) - 1;
#else
            k = __builtin_ctz(low);
#endif
            if (k) {
#pragma unroll
                for (int i = 0; i < N - 1; i++) u[i] = (u[i] >> k) | (u[i + 1] << (32 - k));
                u[N - 1] >>= k;
                // x1 / 2^k mod p
                const uint32_t m = (x1[0] * P::INV) & ((1u << k) - 1u);
                uint64_t c = 0;
                uint32_t t[N + 1];
#pragma unroll
                for (int i = 0; i < N; i++) { c += (uint64_t)m * P::mod(i) + x1[i]; t[i] = (uint32_t)c; c >>= 32; }
                t[N] = (uint32_t)c;
#pragma unroll
                for (int i = 0; i < N; i++) x1[i] = (t[i] >> k) | (t[i + 1] << (32 - k));
            }
        }
        T r;
#pragma unroll
        for (int i = 0; i < N; i++) r.l[i] = x2[i];
        return to_mont(to_mont(r));
    }
};

// Same field, but multiplication is a real (register-ABI) call on the device: the curve kernels issue 10-30
// multiplications per group operation, and inlining them all makes the loop body (35-90 KB of SASS) overflow
// the 32 KB L1.5 instruction cache -- ncu showed "no instruction" as the top stall of k_accumulate.
template <class P>
struct FieldCall : Field<P> {
    typedef typename Field<P>::T T;
    G16_HD_COLD static T mul(T a, T b) { return Field<P>::mul(a, b); }
    G16_HD static T sqr(const T& a) { return mul(a, a); }
};

// Lazy reduction for the hot loops (bucket accumulation): values live in [0, 2p) instead of [0, p), which drops the
// conditional subtraction at the end of every multiplication (17 of its 183 instructions on 8 limbs).  Valid because
// 4p < R for all four moduli (BN254 p, r: 254 bits of 256; BLS12-377 p: 377 of 384, r: 253 of 256): a Montgomery
// product of two values below 2p is below 2p, a sum is below 4p < R and is brought back with one conditional
// subtraction of 2p.  Zero has two representations (0 and p).  canon() returns to [0, p) where values leave the loop.
template <class P>
struct LazyField : Field<P> {
    typedef Field<P> B;
    typedef typename B::T T;
    static constexpr int N = P::N;
    G16_HD static constexpr uint32_t mod2(int i) { return (P::mod(i) << 1) | (i ? P::mod(i - 1) >> 31 : 0u); }
#ifdef __CUDA_ARCH__
    G16_D static T mul(const T& a, const T& b) { return B::mul_raw(a, b); }
    G16_D static T sqr(const T& a) { return mul(a, a); }
    // a < 4p -> [0, 2p)
    G16_D static T reduce2(const T& a) {
        T s;
        s.l[0] = ptx::sub_cc(a.l[0], mod2(0));
#pragma unroll
        for (int i = 1; i < N; i++) s.l[i] = ptx::subc_cc(a.l[i], mod2(i));
        uint32_t borrow = ptx::subc(0, 0);
        T r;
#pragma unroll
        for (int i = 0; i < N; i++) r.l[i] = borrow ? a.l[i] : s.l[i];
        return r;
    }
    G16_D static T add(const T& a, const T& b) {
        T r;
        r.l[0] = ptx::add_cc(a.l[0], b.l[0]);
#pragma unroll
        for (int i = 1; i < N - 1; i++) r.l[i] = ptx::addc_cc(a.l[i], b.l[i]);
        r.l[N - 1] = ptx::addc(a.l[N - 1], b.l[N - 1]);   // a + b < 4p < R
        return reduce2(r);
    }
    G16_D static T dbl(const T& a) { return add(a, a); }
    G16_D static T sub(const T& a, const T& b) {
        T r;
        r.l[0] = ptx::sub_cc(a.l[0], b.l[0]);
#pragma unroll
        for (int i = 1; i < N; i++) r.l[i] = ptx::subc_cc(a.l[i], b.l[i]);
        uint32_t borrow = ptx::subc(0, 0);
        r.l[0] = ptx::add_cc(r.l[0], borrow & mod2(0));
#pragma unroll
        for (int i = 1; i < N - 1; i++) r.l[i] = ptx::addc_cc(r.l[i], borrow & mod2(i));
        r.l[N - 1] = ptx::addc(r.l[N - 1], borrow & mod2(N - 1));
        return r;
    }
    G16_D static bool is_zero(const T& a) {
        uint32_t z = 0, m = 0;
#pragma unroll
        for (int i = 0; i < N; i++) { z |= a.l[i]; m |= a.l[i] ^ P::mod(i); }
        return z == 0 || m == 0;
    }
    G16_D static bool eq(const T& a, const T& b) { return is_zero(sub(a, b)); }
    G16_D static T neg(const T& a) { return B::is_zero(a) ? a : sub(B::zero(), a); }   // 2p - a for a in (0, 2p)
    G16_D static T canon(const T& a) { return B::reduce_once(a); }
#endif
};
// the same with the multiplication as a real call (see FieldCall)
template <class P>
struct LazyFieldCall : LazyField<P> {
    typedef typename Field<P>::T T;
#ifdef __CUDA_ARCH__
    __device__ __noinline__ static T mul(T a, T b) { return Field<P>::mul_raw(a, b); }
    G16_D static T sqr(const T& a) { return mul(a, a); }
#endif
};

#if G16_CURVE == 377
typedef FpBLS377 FpParams;
typedef FrBLS377 FrParams;
constexpr int FP2_NONRESIDUE = 5;    // Fp2 = Fp[u]/(u^2+5)  (gnark-crypto bls12-377 E2)
#else
typedef FpBN254 FpParams;
typedef FrBN254 FrParams;
constexpr int FP2_NONRESIDUE = 1;    // Fp2 = Fp[u]/(u^2+1)  (gnark-crypto bn254 E2)
#endif
typedef Field<FpParams> Fp;
typedef Field<FrParams> Fr;
typedef FieldCall<FpParams> FpC;
typedef LazyFieldCall<FpParams> FpL;
typedef LazyField<FrParams> FrL;

// ---------------------------------------------------------------- Fp2 = Fp[u]/(u^2 + NR)  (gnark E2{A0,A1})
template <class F, int NR>
struct Ext2 {
    struct T { typename F::T a0, a1; };
    G16_HD static T zero() { T r; r.a0 = F::zero(); r.a1 = F::zero(); return r; }
    G16_HD static T one() { T r; r.a0 = F::one(); r.a1 = F::zero(); return r; }
    G16_HD static bool is_zero(const T& a) { return F::is_zero(a.a0) && F::is_zero(a.a1); }
    G16_HD static bool eq(const T& a, const T& b) { return F::eq(a.a0, b.a0) && F::eq(a.a1, b.a1); }
    G16_HD static T add(const T& a, const T& b) { T r; r.a0 = F::add(a.a0, b.a0); r.a1 = F::add(a.a1, b.a1); return r; }
    G16_HD static T sub(const T& a, const T& b) { T r; r.a0 = F::sub(a.a0, b.a0); r.a1 = F::sub(a.a1, b.a1); return r; }
    G16_HD static T dbl(const T& a) { return add(a, a); }
    G16_HD static T neg(const T& a) { T r; r.a0 = F::neg(a.a0); r.a1 = F::neg(a.a1); return r; }
    G16_HD static T canon(const T& a) { T r; r.a0 = F::canon(a.a0); r.a1 = F::canon(a.a1); return r; }
    // NR * a for the small non-residues in use (1 and 5)
    G16_HD static typename F::T mul_nr(const typename F::T& a) {
        static_assert(NR == 1 || NR == 5, "unsupported non-residue");
        if (NR == 1) return a;
        typename F::T a4 = F::dbl(F::dbl(a));
        return F::add(a4, a);
    }
    // Karatsuba: 3 base multiplications
    G16_HD static T mul(const T& a, const T& b) {
        typename F::T v0 = F::mul(a.a0, b.a0), v1 = F::mul(a.a1, b.a1);
        typename F::T s = F::mul(F::add(a.a0, a.a1), F::add(b.a0, b.a1));
        T r;
        r.a0 = F::sub(v0, mul_nr(v1));
        r.a1 = F::sub(F::sub(s, v0), v1);
        return r;
    }
    // 2 base multiplications: (a0+a1)(a0 - NR a1) = (a0^2 - NR a1^2) - (NR-1) a0 a1
    G16_HD static T sqr(const T& a) {
        typename F::T p = F::mul(F::add(a.a0, a.a1), F::sub(a.a0, mul_nr(a.a1)));
        typename F::T q = F::mul(a.a0, a.a1);
        T r;
        if (NR == 1) r.a0 = p;
        else r.a0 = F::add(p, F::dbl(F::dbl(q)));   // NR = 5: + 4 a0 a1
        r.a1 = F::dbl(q);
        return r;
    }
    G16_HD static T inv(const T& a) {
        typename F::T d = F::inv(F::add(F::sqr(a.a0), mul_nr(F::sqr(a.a1))));
        T r; r.a0 = F::mul(a.a0, d); r.a1 = F::neg(F::mul(a.a1, d)); return r;
    }
};
typedef Ext2<FpC, FP2_NONRESIDUE> Fp2;
typedef Ext2<FpL, FP2_NONRESIDUE> Fp2L;     // same memory layout as Fp2, coordinates in [0, 2p)
// Fp2 with the call boundary one level up: mul / sqr are real calls whose 3 / 2 base-field multiplications are inlined
// (a third of the argument shuffling of three separate calls)
template <class F, int NR>
struct Ext2Call : Ext2<F, NR> {
    typedef typename Ext2<F, NR>::T T;
#ifdef __CUDA_ARCH__
    __device__ __noinline__ static T mul(T a, T b) { return Ext2<F, NR>::mul(a, b); }
    __device__ __noinline__ static T sqr(T a) { return Ext2<F, NR>::sqr(a); }
#endif
};
typedef LazyField<FpParams> FpLI;                      // lazily reduced, multiplication inlined
typedef Ext2Call<FpLI, FP2_NONRESIDUE> Fp2LC;

}  // namespace g16
