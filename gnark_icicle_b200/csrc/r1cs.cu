// Sparse matrix-vector products of the R1CS over Fr.  One thread per (matrix, constraint): rows of real circuits
// hold 1-3 terms (a MiMC round: t = x + k + c_i), so a row-per-thread CSR walk is already bandwidth-bound on the
// gathers of W; coefficient one (the common case, gnark's CoeffIdOne) skips the multiplication.
#include "r1cs.cuh"

namespace g16 {

struct R1csArgs {
    const uint32_t* row_ptr[3];
    const uint32_t* wire[3];
    const uint32_t* cid[3];
    Fr::T* out[3];
};

__device__ __forceinline__ Fr::T ld_fr(const Fr::T* p) {
    Fr::T r;
    const uint4* s = reinterpret_cast<const uint4*>(p);
    uint4* d = reinterpret_cast<uint4*>(&r);
    d[0] = __ldg(s); d[1] = __ldg(s + 1);
    return r;
}

// rb, g, rank: this GPU produces the rows whose index bits [rb, rb+g) equal `rank` (the distribution a multi-GPU
// computeH starts from, multi.cu); local row q is global row i.  g = 0: all rows.
__global__ void __launch_bounds__(256) k_r1cs_eval(R1csArgs a, const Fr::T* __restrict__ coeffs, const Fr::T* __restrict__ W,
                                                   uint64_t n_constraints, uint64_t n_local, int rb, int g, uint32_t rank) {
    const uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int m = blockIdx.y;
    if (q >= n_local) return;
    const uint64_t i = ((q >> rb) << (rb + g)) | ((uint64_t)rank << rb) | (q & ((1ull << rb) - 1));
    Fr::T acc = Fr::zero();
    if (i < n_constraints) {
        const uint32_t lo = a.row_ptr[m][i], hi = a.row_ptr[m][i + 1];
        const Fr::T one = Fr::one();
        for (uint32_t t = lo; t < hi; t++) {
            Fr::T w = ld_fr(W + a.wire[m][t]);
            Fr::T cf = ld_fr(coeffs + a.cid[m][t]);
            if (!Fr::eq(cf, one)) w = Fr::mul(w, cf);
            acc = Fr::add(acc, w);
        }
    }
    uint4* d = reinterpret_cast<uint4*>(a.out[m] + q);
    const uint4* s = reinterpret_cast<const uint4*>(&acc);
    d[0] = s[0]; d[1] = s[1];
}

void r1cs_eval(const R1csDev& r, const Fr::T* W_d, Fr::T* a_d, Fr::T* b_d, Fr::T* c_d, uint64_t domain_n, cudaStream_t st,
               int rb, int g, uint32_t rank) {
    domain_n >>= g;   // rows produced here
    if (!domain_n) return;
    R1csArgs a;
    Fr::T* outs[3] = {a_d, b_d, c_d};
    for (int m = 0; m < 3; m++) {
        a.row_ptr[m] = r.row_ptr[m].as<uint32_t>();
        a.wire[m] = r.wire[m].as<uint32_t>();
        a.cid[m] = r.cid[m].as<uint32_t>();
        a.out[m] = outs[m];
    }
    dim3 grid((unsigned)((domain_n + 255) / 256), 3);
    k_r1cs_eval<<<grid, 256, 0, st>>>(a, r.coeffs.as<Fr::T>(), W_d, r.n_constraints, domain_n, rb, g, rank);
    G16_LAUNCH_CHECK();
}

}  // namespace g16
