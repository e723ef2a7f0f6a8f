// Fixed-base batch scalar multiplication [k_i]G on device (setup-side tooling; see msm_api.cuh).
#pragma once
#include "curve_select.h"
#include "msm.cuh"

namespace g16 {

constexpr int FB_WINDOWS = 32;   // 8-bit windows over a 256-bit scalar
constexpr int FB_ROW = 255;

template <class C, class BF>
__global__ void k_fixed_base(const typename C::Aff* __restrict__ table, const Fr::T* __restrict__ scalars, uint64_t n,
                             typename C::Aff* __restrict__ out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr::T s = Fr::from_mont(scalars[i]);
    typename C::Xyzz acc = C::inf();
    for (int j = 0; j < FB_WINDOWS; j++) {
        uint32_t d = (s.l[j >> 2] >> ((j & 3) * 8)) & 0xffu;
        if (d) C::madd(acc, load_aff<C>(table + j * FB_ROW + (d - 1)));
    }
    out[i] = C::to_aff(acc);
}

// host: table[j][d-1] = d * 2^(8j) * G
template <class C>
std::vector<typename C::Aff> fixed_base_table_host(const typename C::Aff& gen) {
    std::vector<typename C::Aff> t((size_t)FB_WINDOWS * FB_ROW);
    typename C::Aff base = gen;
    for (int j = 0; j < FB_WINDOWS; j++) {
        typename C::Xyzz acc = C::from_aff(base);
        t[(size_t)j * FB_ROW] = base;
        for (int d = 2; d <= FB_ROW; d++) {
            C::madd(acc, base);
            t[(size_t)j * FB_ROW + d - 1] = C::to_aff(acc);
        }
        C::madd(acc, base);   // 256 * base
        base = C::to_aff(acc);
    }
    return t;
}

template <class C, class BF>
void fixed_base_run(DevBuf& table_d, const typename C::Aff& gen, const Fr::T* scalars_d, uint64_t n, typename C::Aff* out_d, cudaStream_t st) {
    if (!n) return;
    if (!table_d.p) {
        std::vector<typename C::Aff> t = fixed_base_table_host<C>(gen);
        table_d.reserve(t.size() * sizeof(typename C::Aff));
        G16_CUDA(cudaMemcpyAsync(table_d.p, t.data(), t.size() * sizeof(typename C::Aff), cudaMemcpyHostToDevice, st));
        G16_CUDA(cudaStreamSynchronize(st));
    }
    k_fixed_base<C, BF><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(table_d.as<typename C::Aff>(), scalars_d, n, out_d);
    G16_LAUNCH_CHECK();
}

}  // namespace g16
