// standalone microbenchmark: in-register throughput of Montgomery multiplier variants (8 x 32-bit limbs, BN254 Fp)
//   0: Fp::mul (final conditional subtraction)   1: lazy mul_raw (no final subtraction, values < 2p)
// and of a loop shaped like the XYZZ mixed addition (10 mul + 7 add/sub) in both fields.
#include "../gnark_icicle_b200/csrc/field.cuh"
#include <cstdio>
#include <cuda_runtime.h>
using namespace g16;
template <class F>
__global__ void __launch_bounds__(256) k_mul(uint32_t* sink, uint32_t seed, int iters) {
    typename F::T x[4], y = F::one();
    y.l[0] += seed + threadIdx.x;
    for (int c = 0; c < 4; c++) { x[c] = F::r2(); x[c].l[0] += c + threadIdx.x; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < 4; c++) x[c] = F::mul(x[c], y);
    }
    uint32_t v = 0;
    for (int c = 0; c < 4; c++) for (int k = 0; k < 8; k++) v ^= x[c].l[k];
    if (v == 0x12345678u) sink[0] = v;
}
// madd-shaped: pp = px*zz - x; r = py*zzz - y; p2 = pp^2; p3 = pp*p2; q = x*p2; x3 = r^2 - p3 - 2q; y3 = r(q-x3) - y*p3; zz*=p2; zzz*=p3
template <class F>
__global__ void __launch_bounds__(256, 2) k_madd(uint32_t* sink, uint32_t seed, int iters) {
    typename F::T x = F::r2(), y = F::one(), zz = F::one(), zzz = F::r2(), px = F::one(), py = F::r2();
    x.l[0] += threadIdx.x; y.l[0] += seed; px.l[0] += 2 * threadIdx.x; py.l[0] += 7;
    for (int i = 0; i < iters; i++) {
        typename F::T pp = F::sub(F::mul(px, zz), x), r = F::sub(F::mul(py, zzz), y);
        typename F::T p2 = F::sqr(pp), p3 = F::mul(pp, p2), q = F::mul(x, p2);
        typename F::T x3 = F::sub(F::sub(F::sub(F::sqr(r), p3), q), q);
        y = F::sub(F::mul(r, F::sub(q, x3)), F::mul(y, p3));
        x = x3; zz = F::mul(zz, p2); zzz = F::mul(zzz, p3);
        px.l[0] ^= i; 
    }
    uint32_t v = 0;
    for (int k = 0; k < 8; k++) v ^= x.l[k] ^ y.l[k] ^ zz.l[k] ^ zzz.l[k];
    if (v == 0x12345678u) sink[0] = v;
}
template <class K> float best_of(K launch) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) { cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; }
    return best;
}
int main() {
    uint32_t* sink; cudaMalloc(&sink, 4096); cudaMemset(sink, 0x5a, 4096);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int iters = 512;
    for (int bps = 2; bps <= 8; bps *= 2) {
        int blocks = sms * bps;
        float a = best_of([&] { k_mul<Field<FpBN254>><<<blocks, 256>>>(sink, 3, iters); });
        float b = best_of([&] { k_mul<LazyField<FpBN254>><<<blocks, 256>>>(sink, 3, iters); });
        printf("mul    blocks/SM=%d: canonical %.1f  lazy %.1f G mul/s\n", bps, (double)blocks * 256 * iters * 4 / (a * 1e-3) / 1e9, (double)blocks * 256 * iters * 4 / (b * 1e-3) / 1e9);
    }
    for (int bps = 2; bps <= 4; bps *= 2) {
        int blocks = sms * bps;
        float a = best_of([&] { k_madd<FieldCall<FpBN254>><<<blocks, 256>>>(sink, 3, iters); });
        float b = best_of([&] { k_madd<LazyFieldCall<FpBN254>><<<blocks, 256>>>(sink, 3, iters); });
        float c = best_of([&] { k_madd<Field<FpBN254>><<<blocks, 256>>>(sink, 3, iters); });
        float d = best_of([&] { k_madd<LazyField<FpBN254>><<<blocks, 256>>>(sink, 3, iters); });
        double ops = (double)blocks * 256 * iters * 10;
        printf("madd   blocks/SM=%d: call canonical %.1f  call lazy %.1f  inline canonical %.1f  inline lazy %.1f G mul/s\n", bps,
               ops / (a * 1e-3) / 1e9, ops / (b * 1e-3) / 1e9, ops / (c * 1e-3) / 1e9, ops / (d * 1e-3) / 1e9);
    }
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
}
