// G1 instantiation of the Pippenger pipeline (see msm.cuh).
#include "msm_api.cuh"
#include "fixed_base.cuh"
#include <cstdlib>

namespace g16 {

void msm_g1_enqueue(MsmWorkspace& ws, const Fr::T* scalars_d, const uint32_t* map_d, const G1::Aff* bases_d, uint64_t n,
                    int force_c, int waves, cudaStream_t st, MsmPlan& plan, MsmStats* stats, cudaStream_t tail) {
    static const int variant = [] { const char* e = getenv("G16_ACC_VARIANT_G1"); return e ? atoi(e) : 0; }();
    // 256-thread CTAs, 2 per SM (<= 128 registers): measured 1 % faster than 128 x 4; 128 x 5 spills (profiles/README.md)
    if (variant == 1) msm_enqueue<G1, 128, 4>(ws, scalars_d, map_d, bases_d, n, force_c, waves, st, plan, stats, tail);
    else msm_enqueue<G1, 256, 2>(ws, scalars_d, map_d, bases_d, n, force_c, waves, st, plan, stats, tail);
}
G1::Aff msm_g1_finish(const MsmWorkspace& ws, const MsmPlan& plan) { return msm_finish<G1>(ws, plan); }

G1::Aff g1_generator() {
    G1::Aff g;
    g.x = Fp::one();
    g.y = Fp::add(Fp::one(), Fp::one());
    return g;
}
void fixed_base_g1(DevBuf& table_d, const Fr::T* scalars_d, uint64_t n, G1::Aff* out_d, cudaStream_t st) {
    fixed_base_run<G1, Fp>(table_d, g1_generator(), scalars_d, n, out_d, st);
}

}  // namespace g16
