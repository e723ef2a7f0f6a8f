// does the SMSP overlap the fma (IMAD.WIDE) and alu (IADD3/LOP3/SHF) pipes?  N_ALU alu ops per 8 IMAD.WIDE
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int N_ALU>
__global__ void __launch_bounds__(256) k(uint32_t* sink, uint32_t a0, int iters) {
    uint32_t a = sink[1] + threadIdx.x, b = sink[2] | 1u;
    uint64_t x[8];
    uint32_t y[8];
    for (int i = 0; i < 8; i++) { x[i] = sink[8 + i] + threadIdx.x; y[i] = sink[16 + i] ^ threadIdx.x; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x[i]) : "r"(a), "r"(b));
#pragma unroll
            for (int i = 0; i < N_ALU; i++) asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(y[i & 7]) : "r"(a));
        }
    }
    uint64_t v = 0; for (int i = 0; i < 8; i++) v ^= x[i] ^ y[i];
    if (v == 0x1234567) sink[0] = (uint32_t)v;
}
template <int N_ALU> void run(uint32_t* sink, int sms) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int blocks = sms * 8, iters = 2048; float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0); k<N_ALU><<<blocks, 256>>>(sink, 3, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    double wide = (double)blocks * 256 * iters * 32 / (best * 1e-3) / 1e12;
    printf("alu ops per 8 wide = %2d: IMAD.WIDE %.2f T/s, ALU %.2f T/s\n", N_ALU, wide, wide * N_ALU / 8);
}
int main() {
    uint32_t* sink; cudaMalloc(&sink, 4096); cudaMemset(sink, 0x5a, 4096);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    run<0>(sink, sms); run<2>(sink, sms); run<4>(sink, sms); run<8>(sink, sms); run<12>(sink, sms); run<16>(sink, sms);
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
}
