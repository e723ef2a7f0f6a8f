// Non-template entry points of the MSM translation units (msm_g1.cu, msm_g2.cu).
#pragma once
#include "curve_select.h"
#include "msm.cuh"

namespace g16 {

// pre_c != 0: bases_d is a table of precomputed multiples [255/pre_c + 1][n] made by msm_g1_precompute
void msm_g1_enqueue(MsmWorkspace& ws, const Fr::T* scalars_d, const uint32_t* map_d, const G1::Aff* bases_d, uint64_t n,
                    int force_c, int waves, cudaStream_t st, MsmPlan& plan, MsmStats* stats, cudaStream_t tail = nullptr,
                    int pre_c = 0, MsmWorkspace* sort_from = nullptr, cudaEvent_t acc_gate = nullptr);
G1::Aff msm_g1_finish(const MsmWorkspace& ws, const MsmPlan& plan);
// the same without the final inversion (multi-GPU: the ranks' partial sums are added before going affine)
G1::Xyzz msm_g1_finish_xyzz(const MsmWorkspace& ws, const MsmPlan& plan);
G2::Xyzz msm_g2_finish_xyzz(const MsmWorkspace& ws, const MsmPlan& plan);
void msm_g2_enqueue(MsmWorkspace& ws, const Fr::T* scalars_d, const uint32_t* map_d, const G2::Aff* bases_d, uint64_t n,
                    int force_c, int waves, cudaStream_t st, MsmPlan& plan, MsmStats* stats, cudaStream_t tail = nullptr,
                    int pre_c = 0, MsmWorkspace* sort_from = nullptr, cudaEvent_t acc_gate = nullptr);
// fill rows 1 .. 255/c of table_d (row j = 2^(c j) * row 0); row 0 holds the n bases
void msm_g1_precompute(G1::Aff* table_d, uint64_t n, int c, cudaStream_t st);
void msm_g2_precompute(G2::Aff* table_d, uint64_t n, int c, cudaStream_t st);
G2::Aff msm_g2_finish(const MsmWorkspace& ws, const MsmPlan& plan);

// out[i] = [scalars[i]] * G (generator), affine Montgomery.  Setup-side tooling: the device analogue of
// curve.BatchScalarMultiplicationG1/G2 (backend/groth16/bn254/setup.go:251,320) used to materialise
// synthetic proving keys at BASELINE sizes.  `table_d` caches the window table between calls.
void fixed_base_g1(DevBuf& table_d, const Fr::T* scalars_d, uint64_t n, G1::Aff* out_d, cudaStream_t st);
void fixed_base_g2(DevBuf& table_d, const Fr::T* scalars_d, uint64_t n, G2::Aff* out_d, cudaStream_t st);

G1::Aff g1_generator();
G2::Aff g2_generator();

}  // namespace g16
