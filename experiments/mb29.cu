// standalone microbenchmark: in-register throughput of the 9x29-bit carry-free multiplier vs the 8x32 one
#include "field29.cuh"
#include <cstdio>
#include <cuda_runtime.h>
using namespace g16;
__global__ void __launch_bounds__(256) k_f29(uint32_t* sink, uint32_t seed, int iters) {
    Fp29::T x[4], y;
    for (int k = 0; k < 9; k++) { y.l[k] = (sink[16 + k] + threadIdx.x * 3) & Fp29::MASK; for (int c = 0; c < 4; c++) x[c].l[k] = (sink[32 + c * 9 + k] + threadIdx.x) & Fp29::MASK; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < 4; c++) x[c] = Fp29::mul(x[c], y);
    }
    uint32_t v = 0;
    for (int c = 0; c < 4; c++) for (int k = 0; k < 9; k++) v ^= x[c].l[k];
    if (v == 0x12345678u) sink[0] = v;
}
__global__ void __launch_bounds__(256) k_f32(uint32_t* sink, uint32_t seed, int iters) {
    Fp::T x[4], y = Fp::one();
    y.l[0] += seed + threadIdx.x;
    for (int c = 0; c < 4; c++) { x[c] = Fp::r2(); x[c].l[0] += c + threadIdx.x; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < 4; c++) x[c] = Fp::mul(x[c], y);
    }
    uint32_t v = 0;
    for (int c = 0; c < 4; c++) for (int k = 0; k < 8; k++) v ^= x[c].l[k];
    if (v == 0x12345678u) sink[0] = v;
}
int main() {
    uint32_t* sink; cudaMalloc(&sink, 4096); cudaMemset(sink, 0x5a, 4096);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int which = 0; which < 2; which++) for (int bps = 2; bps <= 8; bps *= 2) {
        int blocks = sms * bps, iters = 512; float best = 1e30f;
        for (int rep = 0; rep < 4; rep++) {
            cudaEventRecord(e0);
            if (which) k_f29<<<blocks, 256>>>(sink, 3, iters); else k_f32<<<blocks, 256>>>(sink, 3, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("%s blocks/SM=%d: %.1f G mul/s\n", which ? "f29" : "f32", bps, (double)blocks * 256 * iters * 4 / (best * 1e-3) / 1e9);
    }
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
}
