// Carry-free Montgomery field for the MSM hot loop: 9 limbs of 29 bits, 64-bit column accumulators.
//
// Why: the 8x32-bit multiplier of field.cuh issues carry-chained IMAD.WIDE.U32.X, which this B200 executes at
// HALF the rate of plain IMAD.WIDE (measured: 8.98 vs 18.41 T/s, profiles/r01_microbench.json), capping it at
// 68 G mul/s.  With 29-bit limbs a 64-bit column absorbs all 18 partial products of a CIOS step without carries,
// so every multiply is a plain full-rate IMAD.WIDE and the carry work moves to the (otherwise idle) ALU pipe.
//
// Representation: value = sum l[i] * 2^(29 i), limbs normalised to [0, 2^29) except the top one; values are kept
// LAZILY reduced in [0, 13p) (capacity 2^261 ~ 169 p).  mul(a, b) = a*b*2^-261 mod p with a < A p, b < B p,
// A*B <= 169 gives a result < 2p.  Elements are therefore in Montgomery form with R' = 2^261; conversion from
// gnark's R = 2^256 image is one multiplication by 2^266 (done once per base at key upload).
// Same role as field.cuh: replaces gnark-crypto fp arithmetic behind G1Jac/G2Jac.MultiExp
// (backend/groth16/bn254/prove.go:204-293) and icicle's device field (icicle.go:302-382).
#pragma once
#include "../gnark_icicle_b200/csrc/field.cuh"

namespace g16 {

struct Fp29BN254 {
    static constexpr uint32_t INV = 0x04866389u;  // -p^-1 mod 2^29
    G16_HD static constexpr uint32_t mod(int i) {
        constexpr uint32_t m[9] = {0x187cfd47u, 0x010460b6u, 0x1c72a34fu, 0x02d522d0u, 0x1585d978u, 0x02db40c0u, 0x00a6e141u, 0x0e5c2634u, 0x0030644eu};
        return m[i];
    }
    // 2^266 mod p and 2^256 mod p as 8x32-bit integers (form conversion constants)
    G16_HD static constexpr uint32_t c266(int i) {
        constexpr uint32_t m[8] = {0x13349ca1u, 0xb34bb095u, 0xf028f972u, 0x1e880124u, 0xa6092b95u, 0xe56cdd25u, 0xdce9ed32u, 0x05800320u};
        return m[i];
    }
    G16_HD static constexpr uint32_t c256(int i) {
        constexpr uint32_t m[8] = {0xc58f0d9du, 0xd35d438du, 0xf5c70b3du, 0x0a78eb28u, 0x7879462cu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return m[i];
    }
};

// a*b + c on 64 bits: one plain (carry-less, full-rate) IMAD.WIDE.U32
G16_HD uint64_t madw(uint32_t a, uint32_t b, uint64_t c) {
#ifdef __CUDA_ARCH__
    uint64_t d;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(d) : "r"(a), "r"(b), "l"(c));
    return d;
#else
    return c + (uint64_t)a * b;
#endif
}
G16_HD uint64_t mulw(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
    uint64_t d;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(d) : "r"(a), "r"(b));
    return d;
#else
    return (uint64_t)a * b;
#endif
}

template <class P>
struct Field29 {
    static constexpr int L = 9;
    static constexpr uint32_t MASK = 0x1fffffffu;
    struct T { uint32_t l[L]; };

    // limb i of K*p, normalised (top limb keeps the excess)
    G16_HD static constexpr uint32_t kp(int K, int i) {
        uint64_t c = 0;
        uint32_t r = 0;
        for (int j = 0; j <= i; j++) {
            c += (uint64_t)K * P::mod(j);
            r = (j < L - 1) ? (uint32_t)(c & MASK) : (uint32_t)c;
            c >>= 29;
        }
        return r;
    }

    G16_HD static T zero() { T r;
#pragma unroll
        for (int i = 0; i < L; i++) r.l[i] = 0; return r; }
    G16_HD static bool is_zero_exact(const T& a) { uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < L; i++) o |= a.l[i]; return o == 0; }

    // 8x32 integer (< 2^256) -> 9x29 limbs
    G16_HD static T from32(const Big<8>& x) {
        T r;
#pragma unroll
        for (int k = 0; k < L; k++) {
            const int bit = 29 * k, w = bit >> 5, s = bit & 31;
            uint64_t two = (uint64_t)x.l[w] | (w + 1 < 8 ? ((uint64_t)x.l[w + 1] << 32) : 0ull);
            r.l[k] = (uint32_t)(two >> s) & MASK;
        }
        return r;
    }
    // normalised 9x29 limbs with value < 2^256 -> 8x32
    G16_HD static Big<8> to32(const T& a) {
        Big<8> r;
        uint64_t acc = 0;
        int nbits = 0, w = 0;
#pragma unroll
        for (int k = 0; k < L; k++) {
            acc |= (uint64_t)a.l[k] << nbits;
            nbits += 29;
            if (nbits >= 32 && w < 8) { r.l[w++] = (uint32_t)acc; acc >>= 32; nbits -= 32; }
        }
        if (w < 8) r.l[w] = (uint32_t)acc;
        return r;
    }

    // propagate carries (signed: limbs may be transiently negative after a subtraction); value must be >= 0
    G16_HD static void normalize(T& a) {
#pragma unroll
        for (int i = 0; i < L - 1; i++) {
            int32_t c = (int32_t)a.l[i] >> 29;
            a.l[i] &= MASK;
            a.l[i + 1] += (uint32_t)c;
        }
    }
    G16_HD static T add(const T& a, const T& b) {
        T r;
#pragma unroll
        for (int i = 0; i < L; i++) r.l[i] = a.l[i] + b.l[i];
        normalize(r);
        return r;
    }
    // a - b + K p  (caller guarantees b < K p, so the value stays non-negative)
    template <int K>
    G16_HD static T sub(const T& a, const T& b) {
        T r;
#pragma unroll
        for (int i = 0; i < L; i++) r.l[i] = a.l[i] - b.l[i] + kp(K, i);
        normalize(r);
        return r;
    }
    // a - b - 2c + K p in one normalisation (x3 of the group law)
    template <int K>
    G16_HD static T sub3(const T& a, const T& b, const T& c) {
        T r;
#pragma unroll
        for (int i = 0; i < L; i++) r.l[i] = a.l[i] - b.l[i] - 2 * c.l[i] + kp(K, i);
        normalize(r);
        return r;
    }
    G16_HD static T dbl(const T& a) { return add(a, a); }

    // a*b*2^-261 mod p, lazily reduced: a < A p, b < B p, A*B <= 169  =>  result < 2p, normalised limbs
    G16_HD static T mul(const T& a, const T& b) {
        uint64_t t[L];
        uint64_t carry = 0;
#pragma unroll
        for (int i = 0; i < L; i++) {
            uint64_t c[L];
            if (i == 0) {
#pragma unroll
                for (int j = 0; j < L; j++) c[j] = mulw(a.l[j], b.l[0]);
            } else {
                c[0] = madw(a.l[0], b.l[i], t[1] + carry);
#pragma unroll
                for (int j = 1; j < L - 1; j++) c[j] = madw(a.l[j], b.l[i], t[j + 1]);
                c[L - 1] = mulw(a.l[L - 1], b.l[i]);
            }
            const uint32_t m = ((uint32_t)c[0] * P::INV) & MASK;
#pragma unroll
            for (int j = 0; j < L; j++) t[j] = madw(m, P::mod(j), c[j]);
            carry = t[0] >> 29;   // the low 29 bits of t[0] are zero by construction
        }
        // result = t >> 29: columns t[1..8]
        T r;
        uint64_t acc = t[1] + carry;
#pragma unroll
        for (int j = 0; j < L - 2; j++) { r.l[j] = (uint32_t)acc & MASK; acc = t[j + 2] + (acc >> 29); }
        r.l[L - 2] = (uint32_t)acc & MASK;
        r.l[L - 1] = (uint32_t)(acc >> 29);
        return r;
    }
    G16_HD static T sqr(const T& a) { return mul(a, a); }

    // x (< K p, normalised) == 0 mod p ?
    template <int K>
    G16_HD static bool is_zero_mod_p(const T& x) {
        bool maybe = false;
#pragma unroll
        for (int k = 0; k < K; k++) maybe |= (x.l[0] == kp(k, 0));
        if (!maybe) return false;
        bool hit = false;
#pragma unroll
        for (int k = 0; k < K; k++) {
            uint32_t d = 0;
#pragma unroll
            for (int i = 0; i < L; i++) d |= x.l[i] ^ kp(k, i);
            hit |= (d == 0);
        }
        return hit;
    }
    // subtract p once if >= p (x < 2p, normalised) -> canonical
    G16_HD static T reduce_once(const T& x) {
        T t;
#pragma unroll
        for (int i = 0; i < L; i++) t.l[i] = x.l[i] - kp(1, i);
#pragma unroll
        for (int i = 0; i < L - 1; i++) {
            int32_t c = (int32_t)t.l[i] >> 29;
            t.l[i] &= MASK;
            t.l[i + 1] += (uint32_t)c;
        }
        return ((int32_t)t.l[L - 1] < 0) ? x : t;
    }
    G16_HD static T k266() { Big<8> b;
#pragma unroll
        for (int i = 0; i < 8; i++) b.l[i] = P::c266(i); return from32(b); }
    G16_HD static T k256() { Big<8> b;
#pragma unroll
        for (int i = 0; i < 8; i++) b.l[i] = P::c256(i); return from32(b); }
    // gnark image (x * 2^256, canonical 8x32) -> lazy R' = 2^261 form
    G16_HD static T from_gnark(const Big<8>& x) { return mul(from32(x), k266()); }
    // lazy R' form (< 13 p) -> gnark image (canonical 8x32, x * 2^256)
    G16_HD static Big<8> to_gnark(const T& x) { return to32(reduce_once(mul(x, k256()))); }
};

typedef Field29<Fp29BN254> Fp29;

}  // namespace g16
