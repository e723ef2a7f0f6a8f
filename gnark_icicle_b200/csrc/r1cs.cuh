// a = L.w, b = R.w, c = O.w on the device: the constraint matrices of a circuit stay resident
// in CSR form, so a proof only uploads the wire vector.  Replaces the a, b, c part of the CPU solver output
// (constraint/bn254/solver.go:540-628) and the three uploads at icicle.go:478-480.
#pragma once
#include "curve_select.h"
#include "common.cuh"
#include "field.cuh"

namespace g16 {

struct R1csDev {
    uint64_t n_constraints = 0, n_wires = 0, n_coeffs = 0;
    DevBuf coeffs;                       // Fr, Montgomery (cs.Coefficients)
    DevBuf row_ptr[3], wire[3], cid[3];  // L, R, O
    void release() {
        coeffs.release();
        for (int m = 0; m < 3; m++) { row_ptr[m].release(); wire[m].release(); cid[m].release(); }
    }
};

// out[m][i] = sum over the terms of row i of matrix m of coeffs[cid] * W[wire], i < n_constraints; rows up to domain_n zeroed
// rb, g, rank != 0: produce only the domain_n >> g rows whose index bits [rb, rb+g) equal rank, packed (multi.cu)
void r1cs_eval(const R1csDev& r, const Fr::T* W_d, Fr::T* a_d, Fr::T* b_d, Fr::T* c_d, uint64_t domain_n, cudaStream_t st,
               int rb = 0, int g = 0, uint32_t rank = 0);

}  // namespace g16
