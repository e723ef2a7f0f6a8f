// G1 instantiation of the Pippenger pipeline (see msm.cuh).
#include "msm_api.cuh"
#include "fixed_base.cuh"
#include <cstdlib>

namespace g16 {

void msm_g1_enqueue(MsmWorkspace& ws, const Fr::T* scalars_d, const uint32_t* map_d, const G1::Aff* bases_d, uint64_t n,
                    int force_c, int waves, cudaStream_t st, MsmPlan& plan, MsmStats* stats, cudaStream_t tail, int pre_c, MsmWorkspace* sort_from, cudaEvent_t acc_gate) {
#if G16_CURVE == 377
    // 12-limb Fp: the accumulator alone is 48 registers; 128 threads x 3 CTAs (<= 168 registers), multiplications as calls
    // (inlined, 10 x 6.4 KB of SASS overflows the instruction cache: 26.6-27.5 vs 24.6 ms -- experiments/README.md)
    msm_enqueue<G1, 128, 3>(ws, scalars_d, map_d, bases_d, n, force_c, waves, st, plan, stats, tail, pre_c, sort_from, acc_gate);
#else
    // 256-thread CTAs, 2 per SM, the field multiplications of the group addition inlined and the loop not unrolled: the
    // body (one XYZZ mixed addition, ~32 KB of SASS) still fits the instruction cache, and the 16 + 8 register moves
    // per multiplication call disappear: 9.68 -> 8.61 ms per 2^22-point launch (63 G Fp-mul/s, 0.93 of the multiplier
    // ceiling).  The rejected launch shapes (call-based 256 x 2 and 128 x 4, inline 128 x 4) are listed in experiments/README.md.
    msm_enqueue<G1, 256, 2, false, 1>(ws, scalars_d, map_d, bases_d, n, force_c, waves, st, plan, stats, tail, pre_c, sort_from, acc_gate);
#endif
}
void msm_g1_precompute(G1::Aff* table_d, uint64_t n, int c, cudaStream_t st) {
    if (!n) return;
    k_precompute<G1><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(table_d, n, c, 255 / c + 1);
    G16_LAUNCH_CHECK();
}
G1::Aff msm_g1_finish(const MsmWorkspace& ws, const MsmPlan& plan) { return msm_finish<G1>(ws, plan); }
G1::Xyzz msm_g1_finish_xyzz(const MsmWorkspace& ws, const MsmPlan& plan) { return msm_finish_xyzz<G1>(ws, plan); }

G1::Aff g1_generator() {
    G1::Aff g;
#if G16_CURVE == 377
    // gnark-crypto bls12-377 G1 generator (checked: on y^2 = x^3 + 1, order r -- tests/test_oracle_bls12_377.py)
    static const uint32_t X[12] = {0xb21be9efu, 0xeab9b16eu, 0xffcd394eu, 0xd5481512u, 0xbd37cb5cu, 0x188282c8u, 0xaa9d41bbu, 0x85951e2cu, 0xbf87ff54u, 0xc8fc6225u, 0xfe740a67u, 0x008848deu};
    static const uint32_t Y[12] = {0x559c8ea6u, 0xfd82de55u, 0x34a9591au, 0xc2fe3d36u, 0x4fb82305u, 0x6d182ad4u, 0xca3e52d9u, 0xbd7fb348u, 0x30afeec4u, 0x1f674f5du, 0xc5102effu, 0x01914a69u};
    Fp::T t;
    memcpy(t.l, X, sizeof(X)); g.x = Fp::to_mont(t);
    memcpy(t.l, Y, sizeof(Y)); g.y = Fp::to_mont(t);
#else
    g.x = Fp::one();
    g.y = Fp::add(Fp::one(), Fp::one());
#endif
    return g;
}
void fixed_base_g1(DevBuf& table_d, const Fr::T* scalars_d, uint64_t n, G1::Aff* out_d, cudaStream_t st) {
    fixed_base_run<G1, Fp>(table_d, g1_generator(), scalars_d, n, out_d, st);
}

}  // namespace g16
