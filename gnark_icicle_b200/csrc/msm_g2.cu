// G2 (Fp2) instantiation of the Pippenger pipeline (see msm.cuh).
#include "msm_api.cuh"
#include "fixed_base.cuh"
#include <cstdlib>

namespace g16 {

void msm_g2_enqueue(MsmWorkspace& ws, const Fr::T* scalars_d, const uint32_t* map_d, const G2::Aff* bases_d, uint64_t n,
                    int force_c, int waves, cudaStream_t st, MsmPlan& plan, MsmStats* stats, cudaStream_t tail) {
    static const int variant = [] { const char* e = getenv("G16_ACC_VARIANT_G2"); return e ? atoi(e) : 0; }();
    // 128 x 2 (194 registers): forcing 3 or 4 CTAs/SM spills and is 30-80 % slower (profiles/README.md)
    (void)variant;
    msm_enqueue<G2, 128, 2>(ws, scalars_d, map_d, bases_d, n, force_c, waves, st, plan, stats, tail);
}
G2::Aff msm_g2_finish(const MsmWorkspace& ws, const MsmPlan& plan) { return msm_finish<G2>(ws, plan); }

// bn254.Generators() G2 (EIP-197), plain integers -> Montgomery
G2::Aff g2_generator() {
    static const uint32_t X0[8] = {0xd992f6edu, 0x46debd5cu, 0xf75edaddu, 0x674322d4u, 0x5e5c4479u, 0x426a0066u, 0x121f1e76u, 0x1800deefu};
    static const uint32_t X1[8] = {0xaef312c2u, 0x97e485b7u, 0x35a9e712u, 0xf1aa4933u, 0x31fb5d25u, 0x7260bfb7u, 0x920d483au, 0x198e9393u};
    static const uint32_t Y0[8] = {0x66fa7daau, 0x4ce6cc01u, 0x0c43d37bu, 0xe3d1e769u, 0x8dcb408fu, 0x4aab7180u, 0xdb8c6debu, 0x12c85ea5u};
    static const uint32_t Y1[8] = {0xd122975bu, 0x55acdadcu, 0x70b38ef3u, 0xbc4b3133u, 0x690c3395u, 0xec9e99adu, 0x585ff075u, 0x090689d0u};
    G2::Aff g;
    Fp::T t;
    memcpy(t.l, X0, 32); g.x.a0 = Fp::to_mont(t);
    memcpy(t.l, X1, 32); g.x.a1 = Fp::to_mont(t);
    memcpy(t.l, Y0, 32); g.y.a0 = Fp::to_mont(t);
    memcpy(t.l, Y1, 32); g.y.a1 = Fp::to_mont(t);
    return g;
}
void fixed_base_g2(DevBuf& table_d, const Fr::T* scalars_d, uint64_t n, G2::Aff* out_d, cudaStream_t st) {
    fixed_base_run<G2, Fp2>(table_d, g2_generator(), scalars_d, n, out_d, st);
}

}  // namespace g16
