// Fr NTT domain for the computeH pipeline (declarations).  See ntt.cu.
#pragma once
#include "curve_select.h"
#include "common.cuh"

namespace g16 {

enum NttTable { TW_FWD = 0, TW_INV = 1, TW_COSET_FWD = 2, TW_COSET_INV = 3 };

// Device-resident twiddles of one power-of-two domain.  Replaces what setupDevicePointers keeps for
// the reference (icicle.go:44-75: CosetTable, CosetTableInv, Den replicated n times, Twiddles,
// TwiddlesInv = 5n field elements): here the coset shift is folded into per-stage twiddles, so a
// transform is a pure butterfly network with no separate scaling or bit-reversal sweep.
//
// Layout of every table: stage with half-size h = 2^s occupies entries [h-1, 2h-1); entry h-1+j is
//   TW_FWD        w_{2h}^j                TW_INV        w_{2h}^-j
//   TW_COSET_FWD  g^(n/2h) * w_{2h}^j     TW_COSET_INV  (g^(n/2h) * w_{2h}^j)^-1
// (n-1 entries per table; built lazily on first use.)
struct NttDomain {
    int logn = 0;
    uint64_t n = 0;
    Fr::T* tw[4] = {nullptr, nullptr, nullptr, nullptr};
    Fr::T n_inv;        // 1/n            (Montgomery)
    Fr::T coset_gen;    // g = 5          (fft.Domain.FrMultiplicativeGen)
    Fr::T coset_inv;    // 1/g
    Fr::T omega;        // primitive n-th root
    Fr::T omega_inv;
    Fr::T h_c1;         // 1/n            -- constants of the fused pointwise step of computeH
    Fr::T h_c2;         // den / n^2,  den = 1/(g^n - 1)   (icicle.go:50-53)
    void init(int logn_);
    const Fr::T* table(NttTable t, cudaStream_t st);
    void release();
};

struct NttBatch {
    Fr::T* p[3];
    int count;
};

// in-place butterfly networks over `count` arrays of n elements
void ntt_dif(NttDomain& d, const NttBatch& b, NttTable t, cudaStream_t st);   // natural in -> bit-reversed out
void ntt_dit(NttDomain& d, const NttBatch& b, NttTable t, cudaStream_t st);   // bit-reversed in -> natural out
// a[i] *= c * base^e(i), e(i) = i (bitrev_index = 0) or bitrev(i)
void fr_scale_powers(Fr::T* a, uint64_t n, int logn, const Fr::T& c, const Fr::T& base, int bitrev_index, int use_base, cudaStream_t st);
// a = (a*b*c1 - c)*c2
void fr_pointwise_h(Fr::T* a, const Fr::T* b, const Fr::T* c, uint64_t n, const Fr::T& c1, const Fr::T& c2, cudaStream_t st);
// computeH on device: a,b,c are n-element device arrays (zero padded); result (bit-reversed order) lands in a
void compute_h_device(NttDomain& d, Fr::T* a, Fr::T* b, Fr::T* c, cudaStream_t st);
void fr_mul_vec(Fr::T* out, const Fr::T* a, const Fr::T* b, uint64_t n, cudaStream_t st);
// computeH with all scratch in a static device array (n <= 2^16): no cudaMalloc / cudaFree; a, b, c (n_constraints
// elements each) are left untouched, h_out receives n elements in bit-reversed order.  See ntt.cu.
void compute_h_noalloc(int device, const Fr::T* a, const Fr::T* b, const Fr::T* c, uint64_t n_constraints, uint64_t n,
                       Fr::T* h_out, cudaStream_t st);

// distributed building blocks (see ntt.cu)
void ntt_dist_stages(NttDomain& d, const NttBatch& b, NttTable t, bool dif, int world_log, uint32_t rank, bool cyclic, int split, cudaStream_t st, int rb = 0);
// fused layout change + peer store between the D(rb) and the block distribution (see ntt.cu); dst[r] = base of rank r's
// receive arrays (count arrays of n_local elements, back to back), peer-mapped
struct DistExchange { Fr::T* dst[8]; const Fr::T* src; int count; uint64_t n_local; int g, L, rb; uint32_t rank; int to_block; };
void ntt_dist_exchange(const DistExchange& a, cudaStream_t st);
void ntt_dist_permute(const Fr::T* src, Fr::T* dst, int count, uint64_t n_local, uint32_t world, int block_layout, int to_wire, cudaStream_t st);

// tile shape of a sweep: log2 of the elements per tile / of the columns per tile
int ntt_log_tile();
int ntt_log_cols();

}  // namespace g16
