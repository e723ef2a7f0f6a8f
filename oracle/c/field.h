/* Montgomery fields on 4x64 limbs (Fr of both curves, BN254 Fp) and 6x64 limbs (BLS12-377 Fp)
 * (oracle; TEST INFRASTRUCTURE ONLY).
 *
 * CPU restatement of the arithmetic the reference obtains from gnark-crypto
 * (github.com/consensys/gnark-crypto v0.14.1-0.20241010154951-6638408a49f3,
 * not vendored; call sites backend/groth16/bn254/prove.go:356-399 for fr,
 * :202-304 for curve ops; backend/groth16/bls12-377/prove.go is the same file over
 * the other curve package).  Element layout = gnark-crypto's: 4 (6) little-endian
 * uint64 limbs, Montgomery form with R = 2^256 (2^384).  ORC_CURVE selects the curve
 * (254 = BN254, the default; 377 = BLS12-377, BASELINE.json config 5).  Parity status: unpinned by
 * golden vectors (see oracle/__init__.py); validated against the Python oracle.
 */
#ifndef ORC_FIELD_H
#define ORC_FIELD_H
#include <stdint.h>
#include <string.h>

typedef unsigned __int128 u128;
#ifndef ORC_CURVE
#define ORC_CURVE 254
#endif

/* 4-limb family: fe, field_t, fe_*, raw_*, f_* */
#define FE fe
#define FT field_t
#define NL 4
#define FEN(x) fe_##x
#define RAW(x) raw_##x
#define FN(x) f_##x
#include "field_tmpl.h"
#undef FE
#undef FT
#undef NL
#undef FEN
#undef RAW
#undef FN
/* 6-limb family: fe6, field6_t, fe6_*, raw6_*, f6_* */
#define FE fe6
#define FT field6_t
#define NL 6
#define FEN(x) fe6_##x
#define RAW(x) raw6_##x
#define FN(x) f6_##x
#include "field_tmpl.h"
#undef FE
#undef FT
#undef NL
#undef FEN
#undef RAW
#undef FN

#if ORC_CURVE == 254
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
#elif ORC_CURVE == 377
/* BLS12-377 moduli (oracle/curve.py: x = 0x8508c00000000001, r = x^4 - x^2 + 1, p = (x-1)^2 r/3 + x) and the
 * Montgomery constants derived from them (checked against Python integers in tests/test_oracle_c_bls12_377.py) */
static const field_t FR = {
    {{0x0a11800000000001ULL, 0x59aa76fed0000001ULL, 0x60b44d1e5c37b001ULL, 0x12ab655e9a2ca556ULL}},
    0x0a117fffffffffffULL,
    {{0x25d577bab861857bULL, 0xcc2c27b58860591fULL, 0xa7cc008fe5dc8593ULL, 0x011fdae7eff1c939ULL}},
    {{0x7d1c7ffffffffff3ULL, 0x7257f50f6ffffff2ULL, 0x16d81575512c0feeULL, 0x0d4bda322bbb9a9dULL}}};
static const field6_t FP = {
    {{0x8508c00000000001ULL, 0x170b5d4430000000ULL, 0x1ef3622fba094800ULL, 0x1a22d9f300f5138fULL, 0xc63b05c06ca1493bULL, 0x01ae3a4617c510eaULL}},
    0x8508bfffffffffffULL,
    {{0xb786686c9400cd22ULL, 0x0329fcaab00431b1ULL, 0x22a5f11162d6b46dULL, 0xbfdf7d03827dc3acULL, 0x837e92f041790bf9ULL, 0x006dfccb1e914b88ULL}},
    {{0x02cdffffffffff68ULL, 0x51409f837fffffb1ULL, 0x9f7db3a98a7d3ff2ULL, 0x7b4e97b76e7c6305ULL, 0x4cf495bf803c84e8ULL, 0x008d6661e2fdf49aULL}}};
#else
#error "ORC_CURVE must be 254 (BN254) or 377 (BLS12-377)"
#endif
#endif
