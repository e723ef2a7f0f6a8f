// G2 (Fp2) instantiation of the Pippenger pipeline (see msm.cuh).
#include "msm_api.cuh"
#include "fixed_base.cuh"
#include <cstdlib>

namespace g16 {

void msm_g2_enqueue(MsmWorkspace& ws, const Fr::T* scalars_d, const uint32_t* map_d, const G2::Aff* bases_d, uint64_t n,
                    int force_c, int waves, cudaStream_t st, MsmPlan& plan, MsmStats* stats, cudaStream_t tail, int pre_c, MsmWorkspace* sort_from, cudaEvent_t acc_gate) {
    // 128 threads x 2 CTAs per SM, base-field multiplications as register-ABI calls.  Measured (profiles/README.md,
    // experiments/README.md): more CTAs per SM do not help (3-4 CTAs/SM: +15..50 %); calling at the Fp2 level (three
    // multiplications inlined per call) needs 252 registers and is 10 % slower; parking the accumulator in shared
    // memory (SmemAcc) costs 15 % on BN254 but removes enough spills on the 12-limb BLS12-377 field to win 5 % there.
#if G16_CURVE == 377
    msm_enqueue<G2, 128, 2, true>(ws, scalars_d, map_d, bases_d, n, force_c, waves, st, plan, stats, tail, pre_c, sort_from, acc_gate);
#else
    msm_enqueue<G2, 128, 2>(ws, scalars_d, map_d, bases_d, n, force_c, waves, st, plan, stats, tail, pre_c, sort_from, acc_gate);
#endif
}
void msm_g2_precompute(G2::Aff* table_d, uint64_t n, int c, cudaStream_t st) {
    if (!n) return;
    k_precompute<G2><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(table_d, n, c, 255 / c + 1);
    G16_LAUNCH_CHECK();
}
G2::Aff msm_g2_finish(const MsmWorkspace& ws, const MsmPlan& plan) { return msm_finish<G2>(ws, plan); }
G2::Xyzz msm_g2_finish_xyzz(const MsmWorkspace& ws, const MsmPlan& plan) { return msm_finish_xyzz<G2>(ws, plan); }

// bn254.Generators() G2 (EIP-197), plain integers -> Montgomery
G2::Aff g2_generator() {
#if G16_CURVE == 377
    // harness generator of the order-r subgroup of the twist y^2 = x^3 + 1/u: cofactor x (the point with x = 2);
    // only the fixed-base test/bench helper uses it (proving-key G2 points come from the caller)
    constexpr int L = 12;
    static const uint32_t X0[L] = {0x9b87bda9u, 0x7faab58cu, 0x8130d2d0u, 0x20a7c4c7u, 0x4cd34f4du, 0x40172476u, 0xf27dce03u, 0x049f123bu, 0x4aee7b99u, 0x35bb0888u, 0x1137c889u, 0x00415abbu};
    static const uint32_t X1[L] = {0x60d27f88u, 0x0ddf884bu, 0x0f56b5e1u, 0x4de60b0cu, 0x9cf3630fu, 0xa208c485u, 0xd42beba1u, 0x24148cd0u, 0x6366d6eau, 0x9d9f07b8u, 0x4ca7d342u, 0x00a272c2u};
    static const uint32_t Y0[L] = {0xbb23ee23u, 0x76521ebbu, 0xb66b5ee5u, 0x33cc6492u, 0x54a524d9u, 0x4ac4ef1au, 0xd09da02fu, 0xf523908eu, 0x237bbb97u, 0xb97bc0e7u, 0xefaacc0du, 0x0199275du};
    static const uint32_t Y1[L] = {0x9c3a8051u, 0xeb2f80acu, 0xf31cbe0cu, 0xe66e6d2du, 0x5ac38999u, 0x78536c11u, 0x7ef09df5u, 0x509504b6u, 0xf85ed167u, 0x4752e212u, 0x68f07eaau, 0x0121cb26u};
#else
    constexpr int L = 8;
    static const uint32_t X0[8] = {0xd992f6edu, 0x46debd5cu, 0xf75edaddu, 0x674322d4u, 0x5e5c4479u, 0x426a0066u, 0x121f1e76u, 0x1800deefu};
    static const uint32_t X1[8] = {0xaef312c2u, 0x97e485b7u, 0x35a9e712u, 0xf1aa4933u, 0x31fb5d25u, 0x7260bfb7u, 0x920d483au, 0x198e9393u};
    static const uint32_t Y0[8] = {0x66fa7daau, 0x4ce6cc01u, 0x0c43d37bu, 0xe3d1e769u, 0x8dcb408fu, 0x4aab7180u, 0xdb8c6debu, 0x12c85ea5u};
    static const uint32_t Y1[8] = {0xd122975bu, 0x55acdadcu, 0x70b38ef3u, 0xbc4b3133u, 0x690c3395u, 0xec9e99adu, 0x585ff075u, 0x090689d0u};
#endif
    G2::Aff g;
    Fp::T t;
    memcpy(t.l, X0, 4 * L); g.x.a0 = Fp::to_mont(t);
    memcpy(t.l, X1, 4 * L); g.x.a1 = Fp::to_mont(t);
    memcpy(t.l, Y0, 4 * L); g.y.a0 = Fp::to_mont(t);
    memcpy(t.l, Y1, 4 * L); g.y.a1 = Fp::to_mont(t);
    return g;
}
void fixed_base_g2(DevBuf& table_d, const Fr::T* scalars_d, uint64_t n, G2::Aff* out_d, cudaStream_t st) {
    fixed_base_run<G2, Fp2>(table_d, g2_generator(), scalars_d, n, out_d, st);
}

}  // namespace g16
