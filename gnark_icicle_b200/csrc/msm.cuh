// Pippenger bucket MSM for sm_100a, templated on the curve (G1 over Fp, G2 over Fp2).
//
// Replaces iciclegnark.MsmOnDevice / MsmG2OnDevice (backend/groth16/bn254/icicle/icicle.go:302,315,332,
// 355,382) and mirrors the results of gnark-crypto's G1Jac/G2Jac.MultiExp as the CPU prover calls them
// (backend/groth16/bn254/prove.go:204,217,237,247,293).  Inputs are gnark memory images: scalars =
// Montgomery Fr, bases = Montgomery affine with (0,0) = infinity (legal, skipped -- the reference must strip
// them on the host, icicle.go:98-105,343-347).  The result is handed back in affine form, the only
// representation that is comparable across implementations.
//
// Pipeline (all on one stream, no host synchronisation until the W window sums are read back):
//   k_digits       from-Montgomery + signed c-bit digit recoding, one biased u16 per (window, scalar)
//   k_hist         per-(window, point-chunk) shared-memory histogram -> global bucket counts
//   k_scan_local/_add  exclusive scan of the W*2^(c-1) counts
//   k_scatter      in-house counting sort: CTA-aggregated range reservation + shared-memory cursors;
//                  emits (point index | sign<<31) grouped by (window, bucket)
//   k_accumulate   balanced segmented reduction: every thread owns S consecutive sorted entries, adds
//                  them with XYZZ mixed additions, and emits whole buckets directly and at most one
//                  head and one tail piece -- long buckets are split across threads/warps automatically
//   k_fixup(+_heavy) stitches head/tail pieces of buckets that span threads (a CTA per long bucket); empty -> infinity
//   k_bucket_reduce / k_window_sum   running-sum reduction sum_b (b+1)*B_b per window, tree-combined
//   host           Horner over the W window sums (W*c doublings) + one inversion to affine
#pragma once
#include "curve_select.h"
#include "common.cuh"
#include <type_traits>

namespace g16 {

struct MsmWorkspace {
    DevBuf digits, sorted, counts, offsets, cursor, buckets, head, tail, partial, wsum, scan_tmp, heavy, pairs, level1, fold;
    // batch-affine pre-reduction (msm_ba.cuh): round outputs (ping-pong), their bucket offsets, per-output descriptors
    // and prefix products, the levels of the inversion tree
    DevBuf ba_pts[2], ba_off[2], ba_desc, ba_prefix, ba_tot[4], ba_pref[3];
    // rounds of batch-affine pair additions before the XYZZ accumulation (0 = none).  Measured on B200, prove at 2^22 with
    // the five MSMs overlapped: 68.1 ms without, 63.3 with 3 rounds on the G2 MSM only, 57.7 with 3 + 3, 56.4 with 4 + 4,
    // 57.0 with 6 + 4 (every round pays a fixed inversion-tree latency, the last ones for little work)
    int ba_rounds = 4;
    bool ba_no_memory = false;   // the round buffers did not fit: stay on the direct path
    PinnedBuf host_wsum;
    cudaEvent_t acc_done = nullptr, tail_done = nullptr, sort_done = nullptr;
    uint64_t sort_n = 0; int sort_c = 0; bool sort_pre = false;   // what `sorted` / `offsets` currently describe
    void release() {
        if (acc_done) cudaEventDestroy(acc_done);
        if (tail_done) cudaEventDestroy(tail_done);
        if (sort_done) cudaEventDestroy(sort_done);
        acc_done = tail_done = sort_done = nullptr;
        digits.release(); sorted.release(); counts.release(); offsets.release(); cursor.release();
        buckets.release(); head.release(); tail.release(); partial.release(); wsum.release(); host_wsum.release(); scan_tmp.release(); heavy.release(); pairs.release(); level1.release(); fold.release();
        for (auto& b : ba_pts) b.release();
        for (auto& b : ba_off) b.release();
        for (auto& b : ba_tot) b.release();
        for (auto& b : ba_pref) b.release();
        ba_desc.release(); ba_prefix.release();
    }
};

struct MsmPlan {
    int c, W;          // window bits; number of bucket SETS (= windows; 1 with precomputed multiples)
    uint32_t B;        // buckets per set = 2^(c-1)
    uint32_t NB;       // W * B
    // bucket reduction: nfold fold levels (k_bucket_fold: fold_m[k] consecutive elements per thread, fold_n[k] elements
    // enter level k, fold_n[0] = B), then k_bucket_reduce over the fold_n[nfold] remaining sums (red_m per thread, red_rb CTAs)
    int nfold, fold_m[2];
    uint32_t fold_n[3];
    int red_m, red_rb;
    int Wd;            // digits (bucket entries) per scalar: = W, or the window count when W == 1 (precomputed mode)
    bool pre;          // bases are a table [Wd][n] of 2^(c j) P_i: every window feeds the same bucket set
};

constexpr int MSM_RED_THREADS = 64;

struct MsmPlan;
inline void msm_plan_reduce(MsmPlan& p);

inline MsmPlan msm_plan(uint64_t n, int force_c, bool g2) {
    MsmPlan p;
    int best = 8; double best_cost = 1e300;
    for (int c = 8; c <= 16; c++) {
        int W = 255 / c + 1;
        double madd = g2 ? 30.0 : 10.0, fadd = g2 ? 42.0 : 14.0;
        // per bucket: fixup + 2 running-sum additions + share of the chunk scalar multiplication, fitted on B200
        double cost = (double)n * W * madd + (double)W * (1u << (c - 1)) * (3.2 * fadd);
        if (cost < best_cost) { best_cost = cost; best = c; }
    }
    p.c = force_c >= 8 && force_c <= 16 ? force_c : best;
    p.W = 255 / p.c + 1;
    p.B = 1u << (p.c - 1);
    p.NB = p.W * p.B;
    msm_plan_reduce(p);
    p.Wd = p.W;
    p.pre = false;
    return p;
}

// Precomputed mode (table of 2^(c j) P_i resident next to the key): one shared bucket set, so a larger window is
// affordable -- the bucket count no longer multiplies with the window count.  Entries per scalar drop from
// ceil(255/16) = 16 to ceil(255/20) = 13 at 2^22 points.
constexpr int MSM_PRE_CMIN = 16, MSM_PRE_CMAX = 21, MSM_PRE_HI_BITS = 8;
inline int msm_pre_choose_c(uint64_t n, bool g2, int forced = 0) {
    if (forced >= MSM_PRE_CMIN && forced <= MSM_PRE_CMAX) return forced;   // g16_ctx_set(ctx, "msm_pre_c", c) before the upload
    // G1, small arrays (the shards of a multi-GPU prove): measured on B200 (experiments/pre_c_sweep.py, resident MSM in ms for
    // c = 16 / 17 / 18 / 19 / 20): 2^19 points 2.23 / 2.27 / 2.70 / 2.48 / 2.56, 2^20: 3.73 / 3.57 / 4.23 / 4.22 / 3.90,
    // 2^21: 6.74 / 6.36 / 7.31 / 7.25 / 6.34 -- the cost model below overrates 18 and 19
#if G16_CURVE != 377   // measured for the 8-limb field only
    if (!g2 && n < 1500000) return n < 750000 ? 16 : 17;
    // G2 pieces (tools/prof_piece.py, with the batch-affine rounds; ms for c = 16 / 17 / 18 / 19 / 20): 730 k points 7.57 / 7.60 / 8.59 /
    // 8.60 / 8.47, 930 k: 8.99 / 8.52 / 9.71 / 9.82 / 9.38, 1.4 M: 11.66 / 11.39 / 12.95 / 12.37 / 11.92, 2.8 M: 19.85 / 19.26 / 21.10 /
    // 20.52 / 18.96 -- the bucket reduction (2 Fp2 XYZZ additions per bucket, a dependent chain of ~100) does not shrink
    // with the piece, the cost model below (per-entry costs from before the rounds) picks 20 for all of them
    if (g2 && n < 2000000) return 17;
#endif
    int best = MSM_PRE_CMIN; double best_cost = 1e300;
    for (int c = MSM_PRE_CMIN; c <= MSM_PRE_CMAX; c++) {
        int W = 255 / c + 1;
        double madd = g2 ? 30.0 : 10.0, fadd = g2 ? 42.0 : 14.0;
        double cost = (double)n * W * madd + (double)(1u << (c - 1)) * (3.2 * fadd);
        if (cost < best_cost) { best_cost = cost; best = c; }
    }
    return best;
}
inline MsmPlan msm_plan_pre(int c) {
    MsmPlan p;
    p.c = c;
    p.W = 1;
    p.Wd = 255 / c + 1;
    p.B = 1u << (c - 1);
    p.NB = p.B;
    msm_plan_reduce(p);
    p.pre = true;
    return p;
}
// Reduction of a bucket set in up to three levels.  A fold level turns n elements into n / m sums (k_bucket_fold, m
// consecutive elements per thread, no scalar multiplication); the last level is the weighted sum of what remains
// (k_bucket_reduce).  Sets of >= 8192 buckets fold until fewer than 8192 sums remain (c = 20: 2^19 -> 2^15 -> 2^12): the
// dependent chain of the last level shrinks from ~86 group additions (one fold of 16, then 16 sums per thread and an
// 11-bit double-and-add) to ~28, and its thread count stays at 2048.  B >= 128 (c >= 8).
inline void msm_plan_reduce(MsmPlan& p) {
    uint32_t cur = p.B;
    p.nfold = 0;
    p.fold_n[0] = cur;
    if (cur >= 8192) {
        // first level by 16 when that still leaves >= 2048 sums (measured at c = 20: 0.38 ms by 16 against 0.48 ms by 8 -- twice
        // the CTAs, each with its own shared-memory tree, and a ragged second wave), the second level by 8
        while (p.nfold < 2 && cur >= 8192) { const uint32_t m = (p.nfold == 0 && cur >= 32768) ? 16 : 8; p.fold_m[p.nfold] = (int)m; cur /= m; p.fold_n[++p.nfold] = cur; }
    } else {
        uint32_t fm = cur / MSM_RED_THREADS;
        if (fm > 16) fm = 16;
        p.fold_m[0] = (int)fm; cur /= fm; p.nfold = 1; p.fold_n[1] = cur;
    }
    uint32_t m = cur / (MSM_RED_THREADS * 32);           // cur is a multiple of 64
    if (m < 1) m = 1;
    p.red_m = (int)m;
    p.red_rb = (int)(cur / (MSM_RED_THREADS * m));
}
inline uint32_t msm_fold_ctas(const MsmPlan& p, int k) { return p.fold_n[k] / (uint32_t)p.fold_m[k] / MSM_RED_THREADS; }

// ------------------------------------------------------------------ digits
static __global__ void k_digits(const Fr::T* __restrict__ scalars, const uint32_t* __restrict__ map, uint64_t n,
                         uint16_t* __restrict__ digits, int c, int W) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t src = map ? map[i] : i;
    const uint4* q = reinterpret_cast<const uint4*>(scalars + src);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    Fr::T s;
    s.l[0] = a.x; s.l[1] = a.y; s.l[2] = a.z; s.l[3] = a.w; s.l[4] = b.x; s.l[5] = b.y; s.l[6] = b.z; s.l[7] = b.w;
    s = Fr::from_mont(s);
    uint32_t l[9];
#pragma unroll
    for (int k = 0; k < 8; k++) l[k] = s.l[k];
    l[8] = 0;
    const uint32_t mask = (1u << c) - 1, half = 1u << (c - 1);
    uint32_t carry = 0;
    for (int w = 0; w < W; w++) {
        int bit = w * c;
        uint32_t raw = 0;
        if (bit < 256) {
            int limb = bit >> 5, sh = bit & 31;
            uint64_t two = (uint64_t)l[limb] | ((uint64_t)l[limb + 1] << 32);
            raw = (uint32_t)(two >> sh) & mask;
        }
        raw += carry;
        // digit d = raw (carry 0) or raw - 2^c (carry 1); stored biased: d + 2^(c-1) in [0, 2^c)
        if (raw >= half) { carry = 1; raw -= half; } else { carry = 0; raw += half; }
        digits[(uint64_t)w * n + i] = (uint16_t)raw;
    }
}

// decode biased digit -> (bucket index |d|-1, sign); returns false for d == 0
__device__ __forceinline__ bool decode_digit(uint32_t enc, uint32_t half, uint32_t& bucket, uint32_t& neg) {
    int d = (int)enc - (int)half;
    if (d == 0) return false;
    neg = d < 0 ? 1u : 0u;
    bucket = (uint32_t)(d < 0 ? -d : d) - 1;
    return true;
}

// ------------------------------------------------------------------ counting sort
static __global__ void k_hist(const uint16_t* __restrict__ digits, uint64_t n, uint32_t* __restrict__ counts, int c, uint64_t per_block) {
    extern __shared__ uint32_t h[];
    const uint32_t B = 1u << (c - 1), half = B;
    const int w = blockIdx.y;
    for (uint32_t b = threadIdx.x; b < B; b += blockDim.x) h[b] = 0;
    __syncthreads();
    uint64_t lo = (uint64_t)blockIdx.x * per_block, hi = lo + per_block;
    if (hi > n) hi = n;
    const uint16_t* dg = digits + (uint64_t)w * n;
    for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        uint32_t bucket, neg;
        if (decode_digit(dg[i], half, bucket, neg)) atomicAdd(&h[bucket], 1u);
    }
    __syncthreads();
    uint32_t* cw = counts + (uint64_t)w * B;
    for (uint32_t b = threadIdx.x; b < B; b += blockDim.x) {
        uint32_t v = h[b];
        if (v) atomicAdd(&cw[b], v);
    }
}

// two-kernel exclusive scan over the W*2^(c-1) bucket counts: CTA-local scan of 4096 entries + CTA totals,
// then every CTA adds the sum of the totals before it.  offsets has nb+1 entries; cursor (nb) gets a copy.
constexpr int SCAN_BLOCK = 4096;
static __global__ void __launch_bounds__(1024) k_scan_local(const uint32_t* __restrict__ counts, uint32_t* __restrict__ offsets,
                                                            uint32_t* __restrict__ block_tot, uint32_t nb) {
    __shared__ uint32_t warp_tot[32];
    const uint32_t base = blockIdx.x * SCAN_BLOCK + threadIdx.x * 4;
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = base + k < nb ? counts[base + k] : 0;
    const uint32_t s = v[0] + v[1] + v[2] + v[3];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = s;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, off); if (lane >= off) inc += t; }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = warp_tot[lane], wi = w;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, wi, off); if (lane >= off) wi += t; }
        warp_tot[lane] = wi - w;
        if (lane == 31) block_tot[blockIdx.x] = wi;
    }
    __syncthreads();
    uint32_t run = inc - s + warp_tot[warp];
#pragma unroll
    for (int k = 0; k < 4; k++) { if (base + k < nb) offsets[base + k] = run; run += v[k]; }
}
static __global__ void __launch_bounds__(1024) k_scan_add(uint32_t* __restrict__ offsets, uint32_t* __restrict__ cursor,
                                                          const uint32_t* __restrict__ block_tot, uint32_t nb) {
    __shared__ uint32_t pre;
    if (threadIdx.x < 32) {
        uint32_t s = 0;
        for (uint32_t i = threadIdx.x; i < blockIdx.x; i += 32) s += block_tot[i];
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
        if (threadIdx.x == 0) pre = s;
    }
    __syncthreads();
    const uint32_t base = blockIdx.x * SCAN_BLOCK + threadIdx.x * 4;
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (base + k < nb) { uint32_t o = offsets[base + k] + pre; offsets[base + k] = o; cursor[base + k] = o; }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) offsets[nb] = pre + block_tot[blockIdx.x];
}

static __global__ void k_scatter(const uint16_t* __restrict__ digits, uint64_t n, uint32_t* __restrict__ cursor,
                          uint32_t* __restrict__ sorted, int c, uint64_t per_block) {
    extern __shared__ uint32_t h[];
    const uint32_t B = 1u << (c - 1), half = B;
    const int w = blockIdx.y;
    for (uint32_t b = threadIdx.x; b < B; b += blockDim.x) h[b] = 0;
    __syncthreads();
    uint64_t lo = (uint64_t)blockIdx.x * per_block, hi = lo + per_block;
    if (hi > n) hi = n;
    const uint16_t* dg = digits + (uint64_t)w * n;
    for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        uint32_t bucket, neg;
        if (decode_digit(dg[i], half, bucket, neg)) atomicAdd(&h[bucket], 1u);
    }
    __syncthreads();
    uint32_t* cw = cursor + (uint64_t)w * B;
    for (uint32_t b = threadIdx.x; b < B; b += blockDim.x) {
        uint32_t v = h[b];
        if (v) h[b] = atomicAdd(&cw[b], v);   // reserve [base, base+v) of this bucket for this CTA
    }
    __syncthreads();
    for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        uint32_t bucket, neg;
        if (decode_digit(dg[i], half, bucket, neg)) {
            uint32_t pos = atomicAdd(&h[bucket], 1u);
            sorted[pos] = (uint32_t)i | (neg << 31);
        }
    }
}

// ------------------------------------------------------------------ precomputed mode: two-level counting sort
// Entries are (bucket, value) pairs, value = sign << 31 | (window * n + point) = an index into the table of
// precomputed multiples.  The 2^(c-1) buckets do not fit a shared-memory histogram, so the sort is MSD in two
// levels: level 1 splits on the top 8 bucket bits (256 groups, CTA-aggregated reservation on 256 global cursors),
// level 2 is the same counting sort as k_hist / k_scatter inside every group on the remaining bits.
template <class F>
__device__ __forceinline__ void for_each_digit(const Fr::T* __restrict__ scalars, const uint32_t* __restrict__ map, uint64_t i,
                                               int c, int W, F&& f) {
    uint64_t src = map ? map[i] : i;
    const uint4* q = reinterpret_cast<const uint4*>(scalars + src);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    Fr::T s;
    s.l[0] = a.x; s.l[1] = a.y; s.l[2] = a.z; s.l[3] = a.w; s.l[4] = b.x; s.l[5] = b.y; s.l[6] = b.z; s.l[7] = b.w;
    s = Fr::from_mont(s);
    uint32_t l[9];
#pragma unroll
    for (int k = 0; k < 8; k++) l[k] = s.l[k];
    l[8] = 0;
    const uint32_t mask = (1u << c) - 1, half = 1u << (c - 1);
    uint32_t carry = 0;
    for (int w = 0; w < W; w++) {
        int bit = w * c;
        uint32_t raw = 0;
        if (bit < 256) {
            int limb = bit >> 5, sh = bit & 31;
            uint64_t two = (uint64_t)l[limb] | ((uint64_t)l[limb + 1] << 32);
            raw = (uint32_t)(two >> sh) & mask;
        }
        raw += carry;
        int d;
        if (raw >= half) { carry = 1; d = (int)raw - (int)(half << 1); } else { carry = 0; d = (int)raw; }
        if (d != 0) f(w, (uint32_t)(d < 0 ? -d : d) - 1, d < 0 ? 1u : 0u);
    }
}

static __global__ void __launch_bounds__(1024) k_pre_hist1(const Fr::T* __restrict__ scalars, const uint32_t* __restrict__ map, uint64_t n,
                                                           int c, int W, int lo_bits, uint32_t* __restrict__ counts1) {
    __shared__ uint32_t h[256];
    if (threadIdx.x < 256) h[threadIdx.x] = 0;
    __syncthreads();
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) for_each_digit(scalars, map, i, c, W, [&](int, uint32_t bucket, uint32_t) { atomicAdd(&h[bucket >> lo_bits], 1u); });
    __syncthreads();
    if (threadIdx.x < 256 && h[threadIdx.x]) atomicAdd(&counts1[threadIdx.x], h[threadIdx.x]);
}
// counts1[256] -> off1[257], cursor1[256]  (one CTA of 256 threads)
static __global__ void __launch_bounds__(256) k_pre_scan1(const uint32_t* __restrict__ counts1, uint32_t* __restrict__ off1, uint32_t* __restrict__ cursor1) {
    __shared__ uint32_t s[256];
    s[threadIdx.x] = counts1[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int k = 0; k < 256; k++) { uint32_t v = s[k]; s[k] = run; run += v; }
        off1[256] = run;
    }
    __syncthreads();
    off1[threadIdx.x] = s[threadIdx.x];
    cursor1[threadIdx.x] = s[threadIdx.x];
}
static __global__ void __launch_bounds__(1024) k_pre_scatter1(const Fr::T* __restrict__ scalars, const uint32_t* __restrict__ map, uint64_t n,
                                                              int c, int W, int lo_bits, uint32_t* __restrict__ cursor1, uint2* __restrict__ pairs) {
    __shared__ uint32_t h[256];
    if (threadIdx.x < 256) h[threadIdx.x] = 0;
    __syncthreads();
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) for_each_digit(scalars, map, i, c, W, [&](int, uint32_t bucket, uint32_t) { atomicAdd(&h[bucket >> lo_bits], 1u); });
    __syncthreads();
    if (threadIdx.x < 256) { uint32_t v = h[threadIdx.x]; if (v) h[threadIdx.x] = atomicAdd(&cursor1[threadIdx.x], v); }
    __syncthreads();
    if (i < n) for_each_digit(scalars, map, i, c, W, [&](int w, uint32_t bucket, uint32_t neg) {
        uint32_t pos = atomicAdd(&h[bucket >> lo_bits], 1u);
        pairs[pos] = make_uint2(bucket, (neg << 31) | (uint32_t)((uint64_t)w * n + i));
    });
}
// grid (chunks, 256 groups): chunk ch of group g covers an equal share of [off1[g], off1[g+1])
__device__ __forceinline__ void pre_chunk(const uint32_t* off1, uint32_t& lo, uint32_t& hi) {
    const uint32_t g0 = off1[blockIdx.y], g1 = off1[blockIdx.y + 1];
    const uint32_t per = (g1 - g0 + gridDim.x - 1) / gridDim.x;
    lo = g0 + blockIdx.x * per;
    hi = lo + per;
    if (lo > g1) lo = g1;
    if (hi > g1) hi = g1;
}
static __global__ void __launch_bounds__(1024) k_pre_hist2(const uint2* __restrict__ pairs, const uint32_t* __restrict__ off1, int lo_bits,
                                                           uint32_t* __restrict__ counts) {
    extern __shared__ uint32_t h[];
    const uint32_t LB = 1u << lo_bits;
    for (uint32_t b = threadIdx.x; b < LB; b += blockDim.x) h[b] = 0;
    __syncthreads();
    uint32_t lo, hi;
    pre_chunk(off1, lo, hi);
    for (uint32_t p = lo + threadIdx.x; p < hi; p += blockDim.x) atomicAdd(&h[pairs[p].x & (LB - 1)], 1u);
    __syncthreads();
    uint32_t* cg = counts + ((uint32_t)blockIdx.y << lo_bits);
    for (uint32_t b = threadIdx.x; b < LB; b += blockDim.x) { uint32_t v = h[b]; if (v) atomicAdd(&cg[b], v); }
}
static __global__ void __launch_bounds__(1024) k_pre_scatter2(const uint2* __restrict__ pairs, const uint32_t* __restrict__ off1, int lo_bits,
                                                              uint32_t* __restrict__ cursor, uint32_t* __restrict__ sorted) {
    extern __shared__ uint32_t h[];
    const uint32_t LB = 1u << lo_bits;
    for (uint32_t b = threadIdx.x; b < LB; b += blockDim.x) h[b] = 0;
    __syncthreads();
    uint32_t lo, hi;
    pre_chunk(off1, lo, hi);
    for (uint32_t p = lo + threadIdx.x; p < hi; p += blockDim.x) atomicAdd(&h[pairs[p].x & (LB - 1)], 1u);
    __syncthreads();
    uint32_t* cg = cursor + ((uint32_t)blockIdx.y << lo_bits);
    for (uint32_t b = threadIdx.x; b < LB; b += blockDim.x) { uint32_t v = h[b]; if (v) h[b] = atomicAdd(&cg[b], v); }
    __syncthreads();
    for (uint32_t p = lo + threadIdx.x; p < hi; p += blockDim.x) {
        uint2 e = pairs[p];
        sorted[atomicAdd(&h[e.x & (LB - 1)], 1u)] = e.y;
    }
}

// table[j * n + i] = 2^(c j) * bases[i], j = 1 .. W-1 (row 0 = the bases themselves, already in place)
template <class C>
__global__ void __launch_bounds__(128) k_precompute(typename C::Aff* __restrict__ table, uint64_t n, int c, int W) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typename C::Aff p = table[i];
    for (int j = 1; j < W; j++) {
        typename C::Xyzz a = C::from_aff(p);
        for (int k = 0; k < c; k++) a = C::dbl(a);
        p = C::to_aff(a);
        table[(uint64_t)j * n + i] = p;
    }
}

// ------------------------------------------------------------------ bucket accumulation
template <class C>
__device__ __forceinline__ typename C::Aff load_aff(const typename C::Aff* p) {
    typename C::Aff r;
    constexpr int Q = sizeof(typename C::Aff) / 16;
    const uint4* src = reinterpret_cast<const uint4*>(p);
    uint4* dst = reinterpret_cast<uint4*>(&r);
#pragma unroll
    for (int i = 0; i < Q; i++) {
        // read-only path with a 64-byte L2 fill hint: the gathers are random, so the default 128-byte line fill of a
        // 64-byte G1 point doubles the DRAM traffic (ncu: 135 B per entry before, profiles/README.md)
        asm volatile("ld.global.nc.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(dst[i].x), "=r"(dst[i].y), "=r"(dst[i].z), "=r"(dst[i].w) : "l"(src + i));
    }
    return r;
}
template <class C>
__device__ __forceinline__ void store_xyzz(typename C::Xyzz* p, const typename C::Xyzz& v) {
    constexpr int Q = sizeof(typename C::Xyzz) / 16;
    const uint4* src = reinterpret_cast<const uint4*>(&v);
    uint4* dst = reinterpret_cast<uint4*>(p);
#pragma unroll
    for (int i = 0; i < Q; i++) dst[i] = src[i];
}
template <class C>
__device__ __forceinline__ typename C::Xyzz load_xyzz(const typename C::Xyzz* p) {
    typename C::Xyzz r;
    constexpr int Q = sizeof(typename C::Xyzz) / 16;
    const uint4* src = reinterpret_cast<const uint4*>(p);
    uint4* dst = reinterpret_cast<uint4*>(&r);
#pragma unroll
    for (int i = 0; i < Q; i++) dst[i] = src[i];
    return r;
}

// entries per thread for `total` sorted entries spread over `nthreads` threads (multiple of 4, >= 4)
__host__ __device__ __forceinline__ uint32_t msm_slice(uint32_t total, uint32_t nthreads) {
    uint32_t s = (total + nthreads - 1) / nthreads;
    s = (s + 3u) & ~3u;
    return s < 4 ? 4 : s;
}

// The running bucket sum of an accumulate thread, in registers (G1) or word-interleaved in shared memory (G2: the
// XYZZ accumulator alone is 64 / 96 registers; parking it in shared memory -- word w of component k of thread t at
// [(k*EW + w) * THREADS + t], conflict-free -- frees enough registers for one more CTA per SM).
// Both run the additions of the loop in the lazily reduced field (values in [0, 2p): no conditional subtraction after
// a multiplication) and bring the sum back to [0, p) when it leaves the thread.
template <class C, int MODE>
struct RegAcc {
    typedef typename LazyCurve<C, MODE>::type CL;
    typename CL::Xyzz a;
    __device__ __forceinline__ void reset() { a = CL::inf(); }
    __device__ __forceinline__ void add(const typename C::Aff& p) { CL::madd(a, reinterpret_cast<const typename CL::Aff&>(p)); }
    __device__ __forceinline__ void store(typename C::Xyzz* dst) const {
        typedef typename CL::BaseField BF;
        typename CL::Xyzz v;
        v.x = BF::canon(a.x); v.y = BF::canon(a.y); v.zz = BF::canon(a.zz); v.zzz = BF::canon(a.zzz);
        store_xyzz<C>(dst, reinterpret_cast<const typename C::Xyzz&>(v));
    }
};
template <class C, int THREADS, int MODE>
struct SmemAcc {
    typedef typename LazyCurve<C, MODE>::type CL;
    typedef typename CL::E E;
    typedef typename CL::BaseField BF;
    static constexpr int EW = sizeof(E) / 4;
    static constexpr size_t BYTES = (size_t)4 * EW * THREADS * 4;
    uint32_t* s;     // shared base + threadIdx.x
    bool is_inf;
    __device__ __forceinline__ E ld(int k) const {
        E r; uint32_t* d = reinterpret_cast<uint32_t*>(&r);
#pragma unroll
        for (int w = 0; w < EW; w++) d[w] = s[(k * EW + w) * THREADS];
        return r;
    }
    __device__ __forceinline__ void st(int k, const E& v) {
        const uint32_t* d = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
        for (int w = 0; w < EW; w++) s[(k * EW + w) * THREADS] = d[w];
    }
    __device__ __forceinline__ void reset() { is_inf = true; }
    // madd-2008-s with every accumulator coordinate loaded right before its use and stored as soon as it is final
    __device__ __forceinline__ void add(const typename C::Aff& p0) {
        const typename CL::Aff& p = reinterpret_cast<const typename CL::Aff&>(p0);
        if (CL::is_inf(p)) return;
        if (is_inf) { st(0, p.x); st(1, p.y); st(2, BF::one()); st(3, BF::one()); is_inf = false; return; }
        E pp = BF::sub(BF::mul(p.x, ld(2)), ld(0));
        E r = BF::sub(BF::mul(p.y, ld(3)), ld(1));
        if (BF::is_zero(pp)) {
            if (BF::is_zero(r)) { typename CL::Xyzz d = CL::dbl_aff(p); st(0, d.x); st(1, d.y); st(2, d.zz); st(3, d.zzz); }
            else is_inf = true;
            return;
        }
        E p2 = BF::sqr(pp);
        E p3 = BF::mul(pp, p2);
        st(2, BF::mul(ld(2), p2));
        E q = BF::mul(ld(0), p2);
        st(3, BF::mul(ld(3), p3));
        E x3 = BF::sub(BF::sub(BF::sub(BF::sqr(r), p3), q), q);
        st(0, x3);
        st(1, BF::sub(BF::mul(r, BF::sub(q, x3)), BF::mul(ld(1), p3)));
    }
    __device__ __forceinline__ void store(typename C::Xyzz* dst) const {
        typename CL::Xyzz v;
        if (is_inf) v = CL::inf();
        else { v.x = BF::canon(ld(0)); v.y = BF::canon(ld(1)); v.zz = BF::canon(ld(2)); v.zzz = BF::canon(ld(3)); }
        store_xyzz<C>(dst, reinterpret_cast<const typename C::Xyzz&>(v));
    }
};

template <class C, int THREADS, int MINB, bool SMEM_ACC, int MODE>
__global__ void __launch_bounds__(THREADS, MINB)
k_accumulate(const typename C::Aff* __restrict__ bases, const uint32_t* __restrict__ sorted,
             const uint32_t* __restrict__ offsets, uint32_t nb, uint32_t nthreads,
             typename C::Xyzz* __restrict__ buckets, typename C::Xyzz* __restrict__ head, typename C::Xyzz* __restrict__ tail) {
    typedef typename C::Xyzz Xyzz;
    typedef typename C::Aff Aff;
    const uint32_t t = blockIdx.x * THREADS + threadIdx.x;
    if (t >= nthreads) return;
    const uint32_t total = offsets[nb];
    const uint32_t S = msm_slice(total, nthreads);
    const uint64_t start64 = (uint64_t)t * S;
    if (start64 >= total) return;
    const uint32_t my_start = (uint32_t)start64;
    const uint32_t my_end = (uint64_t)my_start + S > total ? total : my_start + S;

    // bucket containing my_start: largest gb with offsets[gb] <= my_start and offsets[gb+1] > my_start
    uint32_t lo = 0, hi = nb;  // invariant: offsets[lo] <= my_start < offsets[hi]
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (offsets[mid] <= my_start) lo = mid; else hi = mid;
    }
    uint32_t gb = lo;
    uint32_t b_start = offsets[gb], b_end = offsets[gb + 1];
    uint32_t seg_start = my_start;
    extern __shared__ uint32_t acc_smem[];
    typename std::conditional<SMEM_ACC, SmemAcc<C, THREADS, MODE>, RegAcc<C, MODE>>::type acc;
    if constexpr (SMEM_ACC) acc.s = acc_smem + threadIdx.x;
    acc.reset();

    for (uint32_t p = my_start; p < my_end; p += 4) {
        uint32_t idx[4];
        if (sorted) {
            const uint4 idx4 = *reinterpret_cast<const uint4*>(sorted + p);   // padded allocation: reading past `total` is safe
            idx[0] = idx4.x; idx[1] = idx4.y; idx[2] = idx4.z; idx[3] = idx4.w;
        } else {                                                            // dense, bucket-ordered points (after msm_ba.cuh)
            idx[0] = p; idx[1] = p + 1; idx[2] = p + 2; idx[3] = p + 3;
        }
        // G1 with inlined multiplications: one group addition is ~32 KB of code, so the loop must not be unrolled
#pragma unroll(MODE == 1 && !SMEM_ACC && sizeof(typename C::E) == sizeof(Fp::T) ? 1 : 4)
        for (int k = 0; k < 4; k++) {
            uint32_t pos = p + k;
            if (pos >= my_end) break;
            while (pos == b_end) {   // bucket boundary (also steps over empty buckets)
                if (pos > seg_start) {
                    if (b_start >= my_start) acc.store(buckets + gb);   // bucket lies wholly inside my slice
                    else acc.store(head + t);                           // started in an earlier slice
                    acc.reset();
                    seg_start = pos;
                }
                gb++;
                b_start = b_end;
                b_end = offsets[gb + 1];
            }
            Aff pt = load_aff<C>(bases + (idx[k] & 0x7fffffffu));
            if (idx[k] >> 31) pt = C::neg(pt);
            acc.add(pt);
        }
    }
    // final piece [seg_start, my_end) of bucket gb
    if (my_end > seg_start) {
        bool starts_inside = b_start >= my_start, ends_inside = b_end <= my_end;
        if (starts_inside && ends_inside) acc.store(buckets + gb);
        else if (seg_start == my_start) acc.store(head + t);
        else acc.store(tail + t);
    }
}

// Buckets whose entries span several accumulate slices are stitched here.  Up to FIXUP_SERIAL pieces: one thread
// adds them; longer ones (skewed scalars, the short top window, DummySetup-like inputs) go to a work list that
// k_fixup_heavy reduces with a whole CTA per bucket.
constexpr uint32_t FIXUP_SERIAL = 24;

template <class C>
__global__ void k_fixup(const uint32_t* __restrict__ offsets, uint32_t nb, uint32_t nthreads,
                        typename C::Xyzz* __restrict__ buckets, const typename C::Xyzz* __restrict__ head,
                        const typename C::Xyzz* __restrict__ tail, uint32_t* __restrict__ heavy /* [0] = count, then ids */) {
    typedef typename C::Xyzz Xyzz;
    uint32_t gb = blockIdx.x * blockDim.x + threadIdx.x;
    if (gb >= nb) return;
    const uint32_t total = offsets[nb];
    const uint32_t S = msm_slice(total, nthreads);
    uint32_t start = offsets[gb], end = offsets[gb + 1];
    if (start == end) { store_xyzz<C>(buckets + gb, C::inf()); return; }
    uint32_t t0 = start / S, t1 = (end - 1) / S;
    if (t0 == t1) return;   // written directly by k_accumulate
    if (t1 - t0 + 1 > FIXUP_SERIAL) { heavy[1 + atomicAdd(&heavy[0], 1u)] = gb; return; }
    Xyzz acc = (start == t0 * S) ? load_xyzz<C>(head + t0) : load_xyzz<C>(tail + t0);
    for (uint32_t t = t0 + 1; t <= t1; t++) C::add(acc, load_xyzz<C>(head + t));
    store_xyzz<C>(buckets + gb, acc);
}

template <class C, int THREADS>
__global__ void __launch_bounds__(THREADS)
k_fixup_heavy(const uint32_t* __restrict__ offsets, uint32_t nb, uint32_t nthreads, typename C::Xyzz* __restrict__ buckets,
              const typename C::Xyzz* __restrict__ head, const typename C::Xyzz* __restrict__ tail, const uint32_t* __restrict__ heavy) {
    typedef typename C::Xyzz Xyzz;
    extern __shared__ uint4 fix_sm[];
    Xyzz* sm = reinterpret_cast<Xyzz*>(fix_sm);
    const uint32_t count = heavy[0];
    const uint32_t total = offsets[nb];
    const uint32_t S = msm_slice(total, nthreads);
    for (uint32_t i = blockIdx.x; i < count; i += gridDim.x) {
        const uint32_t gb = heavy[1 + i];
        const uint32_t start = offsets[gb], end = offsets[gb + 1];
        const uint32_t t0 = start / S, t1 = (end - 1) / S;
        Xyzz acc = C::inf();
        for (uint32_t t = t0 + threadIdx.x; t <= t1; t += THREADS) {
            const Xyzz* piece = (t == t0 && start != t0 * S) ? tail + t : head + t;
            C::add(acc, load_xyzz<C>(piece));
        }
        sm[threadIdx.x] = acc;
        __syncthreads();
        for (int stride = THREADS / 2; stride >= 1; stride >>= 1) {
            if ((int)threadIdx.x < stride) { Xyzz a = sm[threadIdx.x]; C::add(a, sm[threadIdx.x + stride]); sm[threadIdx.x] = a; }
            __syncthreads();
        }
        if (threadIdx.x == 0) store_xyzz<C>(buckets + gb, sm[0]);
        __syncthreads();
    }
}

// ------------------------------------------------------------------ bucket reduction
// The weighted sum  sum_k (k+1) B_k  of a bucket set in two levels.  With k = t*m + j (thread t, j < m):
//   (k+1) = (t+1)*m - (m-1-j)   =>   total = m * sum_t (t+1) S_t  -  sum_t T_t,
//   S_t = sum_j B_{tm+j},  T_t = sum_j (m-1-j) B_{tm+j}  -- both from one forward running sum, no scalar multiplication.
// A fold level (k_bucket_fold): S_t -> fold[w][t], T_t tree-summed per CTA -> partialT; the S_t form a bucket set of
// B/m entries that is folded again or goes to k_bucket_reduce.  The host finishes with log2(m) doublings and one
// subtraction per level and set.
// grid (B / m / 64, W)
template <class C>
__global__ void __launch_bounds__(MSM_RED_THREADS)
k_bucket_fold(const typename C::Xyzz* __restrict__ buckets, uint32_t B, int m, typename C::Xyzz* __restrict__ fold,
              typename C::Xyzz* __restrict__ partialT) {
    typedef typename C::Xyzz Xyzz;
    __shared__ Xyzz sm[MSM_RED_THREADS];
    const int w = blockIdx.y;
    const uint32_t t = blockIdx.x * MSM_RED_THREADS + threadIdx.x;
    const Xyzz* bw = buckets + (uint64_t)w * B + (uint64_t)t * m;
    Xyzz run = C::inf(), ws = C::inf();
    for (int j = 0; j < m; j++) {
        C::add(ws, run);
        C::add(run, load_xyzz<C>(bw + j));
    }
    store_xyzz<C>(fold + (uint64_t)w * (B / m) + t, run);
    sm[threadIdx.x] = ws;
    __syncthreads();
    for (int stride = MSM_RED_THREADS / 2; stride >= 1; stride >>= 1) {
        if ((int)threadIdx.x < stride) { Xyzz a = sm[threadIdx.x]; C::add(a, sm[threadIdx.x + stride]); sm[threadIdx.x] = a; }
        __syncthreads();
    }
    if (threadIdx.x == 0) store_xyzz<C>(partialT + (uint64_t)w * gridDim.x + blockIdx.x, sm[0]);
}

// grid (red_rb, W); thread handles m consecutive buckets; CTA result -> partial[w * red_rb + blockIdx.x]
template <class C>
__global__ void __launch_bounds__(MSM_RED_THREADS)
k_bucket_reduce(const typename C::Xyzz* __restrict__ buckets, uint32_t B, int m, typename C::Xyzz* __restrict__ partial) {
    typedef typename C::Xyzz Xyzz;
    __shared__ Xyzz sm[MSM_RED_THREADS];
    const int w = blockIdx.y;
    const uint32_t chunk = blockIdx.x * MSM_RED_THREADS + threadIdx.x;
    const uint32_t b0 = chunk * (uint32_t)m;
    const Xyzz* bw = buckets + (uint64_t)w * B;
    Xyzz run = C::inf(), ws = C::inf();
    for (int k = m - 1; k >= 0; k--) {
        C::add(run, load_xyzz<C>(bw + b0 + k));
        C::add(ws, run);
    }
    // + b0 * run  (weights are b+1; the running sums above cover (b - b0 + 1))
    if (b0) {   // double-and-add from the top set bit of b0
        Xyzz t = run;
        for (int i = 30 - __clz(b0); i >= 0; i--) {
            t = C::dbl(t);
            if ((b0 >> i) & 1) C::add(t, run);
        }
        C::add(ws, t);
    }
    sm[threadIdx.x] = ws;
    __syncthreads();
    for (int stride = MSM_RED_THREADS / 2; stride >= 1; stride >>= 1) {
        if ((int)threadIdx.x < stride) { Xyzz a = sm[threadIdx.x]; C::add(a, sm[threadIdx.x + stride]); sm[threadIdx.x] = a; }
        __syncthreads();
    }
    if (threadIdx.x == 0) store_xyzz<C>(partial + (uint64_t)w * gridDim.x + blockIdx.x, sm[0]);
}

// grid (W, segments), 128 threads: segment y sums its cnt[y] partials of window x (partial + off[y] + x * cnt[y])
constexpr int MSM_WSUM_THREADS = 128;
struct WsumArgs { uint32_t off[3], cnt[3]; };
template <class C>
__global__ void __launch_bounds__(MSM_WSUM_THREADS) k_window_sum(const typename C::Xyzz* __restrict__ partial, WsumArgs a, typename C::Xyzz* __restrict__ wsum) {
    typedef typename C::Xyzz Xyzz;
    extern __shared__ uint4 wsum_sm[];
    Xyzz* sm = reinterpret_cast<Xyzz*>(wsum_sm);
    const int w = blockIdx.x, seg = blockIdx.y;
    const uint32_t rb = a.cnt[seg];
    const Xyzz* src = partial + a.off[seg] + (uint64_t)w * rb;
    Xyzz mine = C::inf();
    for (uint32_t i = threadIdx.x; i < rb; i += MSM_WSUM_THREADS) C::add(mine, load_xyzz<C>(src + i));
    sm[threadIdx.x] = mine;
    __syncthreads();
    for (int stride = MSM_WSUM_THREADS / 2; stride >= 1; stride >>= 1) {
        if ((int)threadIdx.x < stride) { Xyzz x = sm[threadIdx.x]; C::add(x, sm[threadIdx.x + stride]); sm[threadIdx.x] = x; }
        __syncthreads();
    }
    if (threadIdx.x == 0) store_xyzz<C>(wsum + (uint64_t)seg * gridDim.x + w, sm[0]);
}

}  // namespace g16
#include "msm_ba.cuh"
namespace g16 {

// ------------------------------------------------------------------ host driver
struct MsmStats { int c = 0, W = 0; uint32_t acc_threads = 0; cudaEvent_t acc_begin = nullptr, acc_end = nullptr; };

// Enqueues the whole MSM on `st`; window sums land in ws.host_wsum (pinned) after the stream drains.
template <class C, int ACC_THREADS, int ACC_MINB, bool SMEM_ACC = false, int MODE = 0>
void msm_enqueue(MsmWorkspace& ws, const Fr::T* scalars_d, const uint32_t* map_d, const typename C::Aff* bases_d,
                 uint64_t n, int force_c, int waves, cudaStream_t st, MsmPlan& plan_out, MsmStats* stats,
                 cudaStream_t tail = nullptr, int pre_c = 0, MsmWorkspace* sort_from = nullptr, cudaEvent_t acc_gate = nullptr) {
    // acc_gate: the accumulation (long-lived CTAs that fill every SM) is not launched before this event; the sort is
    // pre_c != 0: bases_d is the table [255/pre_c + 1][n] of precomputed multiples (k_precompute)
    // sort_from: another MSM over the SAME scalars, map and n (bs1 / bs2 of a proof) whose sorted entries and bucket
    // offsets are reused when the plans agree -- its stream must already have the sort enqueued
    typedef typename C::Xyzz Xyzz;
    constexpr bool IS_G2 = sizeof(typename C::E) > sizeof(Fp::T);   // the base field is Fp2
    MsmPlan pl = pre_c ? msm_plan_pre(pre_c) : msm_plan(n, force_c, IS_G2);
    plan_out = pl;
    if (stats) { stats->c = pl.c; stats->W = pl.Wd; stats->acc_threads = 0; }
    if (n == 0) return;   // callers treat an empty MSM as infinity (icicle.go:330-333)
    if ((uint64_t)n * pl.Wd >= 0x7ffffff0ull || n >= 0x7fffffffull) throw std::runtime_error("MSM too large for 32-bit entry indices");
    const uint32_t total_max = (uint32_t)(n * pl.Wd);

    int dev = 0, sms = 0;
    G16_CUDA(cudaGetDevice(&dev));
    G16_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int occ = 0;
    auto acc_kernel = k_accumulate<C, ACC_THREADS, ACC_MINB, SMEM_ACC, MODE>;
    const size_t acc_smem = SMEM_ACC ? SmemAcc<C, ACC_THREADS, MODE>::BYTES : 0;
    if (acc_smem > 48 * 1024) G16_CUDA(cudaFuncSetAttribute(acc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)acc_smem));
    G16_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, acc_kernel, ACC_THREADS, acc_smem));
    if (occ < 1) occ = 1;
    uint32_t nthreads = (uint32_t)sms * occ * ACC_THREADS * (waves > 0 ? waves : 2);
    // at least 16 entries per thread: shorter slices only multiply the pieces k_fixup has to stitch (a bucket of the
    // shared set holds ~32 entries at 2^16 points, and every bucket cut into > 8 pieces goes through k_fixup_heavy)
    uint32_t cap = (total_max + 15) / 16;
    if (nthreads > cap) nthreads = cap < 1 ? 1 : cap;

    if (!pl.pre) ws.digits.reserve((size_t)n * pl.W * sizeof(uint16_t) + 64);
    else { ws.pairs.reserve((size_t)total_max * sizeof(uint2) + 64); ws.level1.reserve(3 * 260 * sizeof(uint32_t)); }
    ws.sorted.reserve((size_t)total_max * sizeof(uint32_t) + 64);
    ws.counts.reserve((size_t)pl.NB * 4);
    ws.offsets.reserve((size_t)(pl.NB + 1) * 4);
    ws.cursor.reserve((size_t)pl.NB * 4);
    ws.heavy.reserve((size_t)(pl.NB + 1) * 4);
    ws.scan_tmp.reserve((size_t)((pl.NB + SCAN_BLOCK - 1) / SCAN_BLOCK) * 4 + 16);
    ws.buckets.reserve((size_t)pl.NB * sizeof(Xyzz));
    ws.head.reserve((size_t)nthreads * sizeof(Xyzz));
    ws.tail.reserve((size_t)nthreads * sizeof(Xyzz));
    {
        size_t fold_elems = 0, partials = pl.red_rb;
        for (int k = 0; k < pl.nfold; k++) { fold_elems += pl.fold_n[k + 1]; partials += msm_fold_ctas(pl, k); }
        ws.fold.reserve((size_t)pl.W * fold_elems * sizeof(Xyzz));
        ws.partial.reserve((size_t)pl.W * partials * sizeof(Xyzz));
    }
    ws.wsum.reserve((size_t)3 * pl.W * sizeof(Xyzz));
    ws.host_wsum.reserve((size_t)3 * pl.W * sizeof(Xyzz));

    uint32_t* sorted = ws.sorted.as<uint32_t>();
    uint32_t* counts = ws.counts.as<uint32_t>();
    uint32_t* offsets = ws.offsets.as<uint32_t>();
    uint32_t* cursor = ws.cursor.as<uint32_t>();
    Xyzz* buckets = ws.buckets.as<Xyzz>();
    const uint32_t scan_blocks = (pl.NB + SCAN_BLOCK - 1) / SCAN_BLOCK;
    if (sort_from && sort_from->sort_done && sort_from->sort_n == n && sort_from->sort_c == pl.c && sort_from->sort_pre == pl.pre) {
        G16_CUDA(cudaStreamWaitEvent(st, sort_from->sort_done, 0));
        sorted = sort_from->sorted.as<uint32_t>();
        offsets = sort_from->offsets.as<uint32_t>();
    } else if (pl.pre) {
        const int lo_bits = pl.c - 1 - MSM_PRE_HI_BITS;
        uint32_t* counts1 = ws.level1.as<uint32_t>();
        uint32_t* off1 = counts1 + 260;
        uint32_t* cursor1 = off1 + 260;
        uint2* pairs = ws.pairs.as<uint2>();
        const unsigned sblocks = (unsigned)((n + 1023) / 1024);
        G16_CUDA(cudaMemsetAsync(counts1, 0, 256 * 4, st));
        G16_CUDA(cudaMemsetAsync(counts, 0, (size_t)pl.NB * 4, st));
        k_pre_hist1<<<sblocks, 1024, 0, st>>>(scalars_d, map_d, n, pl.c, pl.Wd, lo_bits, counts1);
        G16_LAUNCH_CHECK();
        k_pre_scan1<<<1, 256, 0, st>>>(counts1, off1, cursor1);
        G16_LAUNCH_CHECK();
        k_pre_scatter1<<<sblocks, 1024, 0, st>>>(scalars_d, map_d, n, pl.c, pl.Wd, lo_bits, cursor1, pairs);
        G16_LAUNCH_CHECK();
        // level 2: enough chunks per group to fill the machine a few times, at least ~8K entries each
        uint32_t chunks2 = (uint32_t)((sms * 4 + 255) / 256);
        uint64_t per_group = ((uint64_t)total_max + 255) / 256;
        while (chunks2 > 1 && per_group / chunks2 < 8192) chunks2--;
        dim3 g2grid(chunks2, 256);
        const size_t sm2 = (size_t)4 << lo_bits;
        k_pre_hist2<<<g2grid, 1024, sm2, st>>>(pairs, off1, lo_bits, counts);
        G16_LAUNCH_CHECK();
        k_scan_local<<<scan_blocks, 1024, 0, st>>>(counts, offsets, ws.scan_tmp.as<uint32_t>(), pl.NB);
        G16_LAUNCH_CHECK();
        k_scan_add<<<scan_blocks, 1024, 0, st>>>(offsets, cursor, ws.scan_tmp.as<uint32_t>(), pl.NB);
        G16_LAUNCH_CHECK();
        k_pre_scatter2<<<g2grid, 1024, sm2, st>>>(pairs, off1, lo_bits, cursor, sorted);
        G16_LAUNCH_CHECK();
    } else {
    uint16_t* digits = ws.digits.as<uint16_t>();

    k_digits<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(scalars_d, map_d, n, digits, pl.c, pl.W);
    G16_LAUNCH_CHECK();
    G16_CUDA(cudaMemsetAsync(counts, 0, (size_t)pl.NB * 4, st));

    // (window, point-chunk) CTAs; enough chunks to fill the machine a few times over
    size_t hsmem = (size_t)pl.B * 4;
    uint32_t chunks = (uint32_t)((sms * 4 + pl.W - 1) / pl.W);
    uint64_t per_block = (n + chunks - 1) / chunks;
    if (per_block < 1024) per_block = 1024;
    chunks = (uint32_t)((n + per_block - 1) / per_block);
    if (hsmem > 48 * 1024) {
        G16_CUDA(cudaFuncSetAttribute(k_hist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem));
        G16_CUDA(cudaFuncSetAttribute(k_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem));
    }
    dim3 hgrid(chunks, pl.W);
    k_hist<<<hgrid, 1024, hsmem, st>>>(digits, n, counts, pl.c, per_block);
    G16_LAUNCH_CHECK();
    k_scan_local<<<scan_blocks, 1024, 0, st>>>(counts, offsets, ws.scan_tmp.as<uint32_t>(), pl.NB);
    G16_LAUNCH_CHECK();
    k_scan_add<<<scan_blocks, 1024, 0, st>>>(offsets, cursor, ws.scan_tmp.as<uint32_t>(), pl.NB);
    G16_LAUNCH_CHECK();
    k_scatter<<<hgrid, 1024, hsmem, st>>>(digits, n, cursor, sorted, pl.c, per_block);
    G16_LAUNCH_CHECK();
    }

    if (sorted == ws.sorted.as<uint32_t>()) {
        if (!ws.sort_done) G16_CUDA(cudaEventCreateWithFlags(&ws.sort_done, cudaEventDisableTiming));
        ws.sort_n = n; ws.sort_c = pl.c; ws.sort_pre = pl.pre;
        G16_CUDA(cudaEventRecord(ws.sort_done, st));
    }
    // ---- batch-affine rounds: every round halves the entries of every bucket with 6M instead of 10M per addition
    const typename C::Aff* acc_bases = bases_d;
    uint32_t acc_total_max = total_max;
    // every round pays a fixed inversion-tree latency (~0.4-0.6 ms, one Fermat chain): small MSMs -- the pieces of a
    // multi-GPU prove, which have little else running beside them -- get fewer rounds
    int ba_rounds = (ws.ba_rounds > 0 && !ws.ba_no_memory && n >= (1u << 17)) ? (ws.ba_rounds > 6 ? 6 : ws.ba_rounds) : 0;
    static const bool ba_noclamp = getenv("G16_BA_NOCLAMP") != nullptr;   // experiments (tools/prof_piece.py --ba): take "msm_ba_rounds" as given
    if (!ba_noclamp) {
        // measured alone on a B200 (tools/prof_piece.py --ba 1 2 3 4 5 with G16_BA_NOCLAMP=1; ms for 1 / 2 / 3 / 4 rounds).  G1: 2^18 points
        // 1.73 / 1.85 / 2.04 / 2.29, 2^19: 2.38 / 2.61 / 2.73 / 2.93, 2^20: 3.78 / 3.88 / 4.04 / 4.19, 2.7 M: 7.58 / 7.48 / 7.49 / 7.79,
        // 4.2 M: 11.19 / 10.86 / 10.78 / 10.99.  G2: 0.73 M 7.86 / 7.32 / 7.33 / 7.38, 1.4 M: 12.69 / 11.39 / 11.12 / 11.11, 2.8 M:
        // 21.54 / 19.44 / 18.49 / 18.60.  With five MSMs in flight the multiplication count matters more than the tree latency,
        // which is why the large arrays of a whole key keep 4 rounds (prove at 2^22: 57.7 ms with 3 + 3, 56.4 with 4 + 4).
        if (IS_G2) {
            if (ba_rounds > 3 && total_max < (32u << 20)) ba_rounds = 3;
            if (ba_rounds > 2 && total_max < (6u << 20)) ba_rounds = 2;
        } else {
            if (ba_rounds > 3 && total_max < (32u << 20)) ba_rounds = 3;
            if (ba_rounds > 2 && total_max < (20u << 20)) ba_rounds = 2;
            if (ba_rounds > 1 && total_max < (12u << 20)) ba_rounds = 1;
        }
    }
    // the gate exists for the long-lived CTAs of a direct XYZZ accumulation (~180 entries per thread); the kernels of the
    // batch-affine rounds are short-lived, and after >= 3 rounds so is the accumulation that follows them
    if (acc_gate && ba_rounds < 3) G16_CUDA(cudaStreamWaitEvent(st, acc_gate, 0));
    if (stats && stats->acc_begin) G16_CUDA(cudaEventRecord(stats->acc_begin, st));
    if (ba_rounds) {
        // the same groups over the lazily reduced fields (values in [0, 2p) between multiplications; G1 with the
        // multiplications inlined): identical memory layout, canonical values in memory
        typedef typename LazyCurve<C, (IS_G2 || G16_CURVE == 377) ? 0 : 1>::type CB;   // 12-limb bodies would overflow the instruction cache
        typedef typename CB::E E;
        typedef typename CB::Aff Aff;
        typedef BaNorm<typename CB::BaseField> BN;      // the shared inversion runs in the prime field (norms of Fp2 denominators)
        typedef typename BN::NF NF;
        typedef typename NF::T NE;
        uint64_t out_max[8], in_max = total_max;
        for (int r = 0; r < ba_rounds; r++) { out_max[r] = in_max / 2 + pl.NB; in_max = out_max[r]; }
        const uint64_t t0_max = (out_max[0] + 32 * BA_K - 1) / (32 * BA_K) * 32;
        std::vector<uint64_t> lvl{t0_max};
        while (lvl.back() > 4096 && lvl.size() < 4) lvl.push_back((lvl.back() + BA_G - 1) / BA_G);
        size_t need = out_max[0] * (sizeof(Aff) + 4 + sizeof(E)) + (ba_rounds > 1 ? out_max[1] * sizeof(Aff) : 0) + 2 * t0_max * sizeof(E) + 2 * (size_t)(pl.NB + 1) * 4;
        if (ws.ba_pts[0].cap < out_max[0] * sizeof(Aff)) {          // first use at this size: only if it fits comfortably
            size_t free_b = 0, total_b = 0;
            G16_CUDA(cudaMemGetInfo(&free_b, &total_b));
            if (need > free_b / 5 * 3) ws.ba_no_memory = true;
        }
        if (!ws.ba_no_memory) {
            ws.ba_pts[0].reserve(out_max[0] * sizeof(Aff) + 64);
            if (ba_rounds > 1) ws.ba_pts[1].reserve(out_max[1] * sizeof(Aff) + 64);
            ws.ba_off[0].reserve((size_t)(pl.NB + 1) * 4);
            ws.ba_off[1].reserve((size_t)(pl.NB + 1) * 4);
            ws.ba_desc.reserve(out_max[0] * 4);
            ws.ba_prefix.reserve(out_max[0] * sizeof(E));
            for (size_t l = 0; l < lvl.size(); l++) ws.ba_tot[l].reserve(lvl[l] * sizeof(E));
            for (size_t l = 0; l + 1 < lvl.size(); l++) ws.ba_pref[l].reserve(lvl[l] * sizeof(E));
            BaSrc<CB> src{reinterpret_cast<const Aff*>(bases_d), sorted};
            const uint32_t* off_in = offsets;
            for (int r = 0; r < ba_rounds; r++) {
                uint32_t* off_out = ws.ba_off[r & 1].as<uint32_t>();
                Aff* out = ws.ba_pts[r & 1].as<Aff>();
                k_ba_counts<<<(pl.NB + 255) / 256, 256, 0, st>>>(off_in, pl.NB, counts);
                G16_LAUNCH_CHECK();
                k_scan_local<<<scan_blocks, 1024, 0, st>>>(counts, off_out, ws.scan_tmp.as<uint32_t>(), pl.NB);
                G16_LAUNCH_CHECK();
                k_scan_add<<<scan_blocks, 1024, 0, st>>>(off_out, cursor, ws.scan_tmp.as<uint32_t>(), pl.NB);
                G16_LAUNCH_CHECK();
                const uint32_t nthr = (uint32_t)((out_max[r] + 32 * BA_K - 1) / (32 * BA_K)) * 32;   // whole warps of 32 K outputs
                const unsigned grid = (nthr + BA_THREADS - 1) / BA_THREADS;
                NE* tot0 = ws.ba_tot[0].as<NE>();
                NE* prefix = ws.ba_prefix.as<NE>();
                NE* nrm = BN::EXT ? prefix + out_max[0] : nullptr;      // Fp2: the second half of the prefix buffer holds the norms
                k_ba_fwd<CB><<<grid, BA_THREADS, 0, st>>>(src, off_in, off_out, pl.NB, ws.ba_desc.as<uint32_t>(), prefix, nrm, tot0, nthr);
                G16_LAUNCH_CHECK();
                // inversion tree over the nthr per-thread totals
                std::vector<uint32_t> sz{nthr};
                while (sz.back() > 4096 && sz.size() < 4) sz.push_back((sz.back() + BA_G - 1) / BA_G);
                for (size_t l = 0; l + 1 < sz.size(); l++) {
                    k_ba_up<NF><<<(sz[l + 1] + BA_THREADS - 1) / BA_THREADS, BA_THREADS, 0, st>>>(ws.ba_tot[l].as<NE>(), sz[l], ws.ba_pref[l].as<NE>(), ws.ba_tot[l + 1].as<NE>(), sz[l + 1]);
                    G16_LAUNCH_CHECK();
                }
                k_ba_leaf<NF><<<(sz.back() + BA_THREADS - 1) / BA_THREADS, BA_THREADS, 0, st>>>(ws.ba_tot[sz.size() - 1].as<NE>(), sz.back());
                G16_LAUNCH_CHECK();
                for (size_t l = sz.size() - 1; l-- > 0;) {
                    k_ba_down<NF><<<(sz[l + 1] + BA_THREADS - 1) / BA_THREADS, BA_THREADS, 0, st>>>(ws.ba_tot[l].as<NE>(), sz[l], ws.ba_pref[l].as<NE>(), ws.ba_tot[l + 1].as<NE>(), ws.ba_tot[l].as<NE>(), sz[l + 1]);
                    G16_LAUNCH_CHECK();
                }
                k_ba_bwd<CB><<<grid, BA_THREADS, 0, st>>>(src, off_out, pl.NB, ws.ba_desc.as<uint32_t>(), prefix, nrm, tot0, out, nthr);
                G16_LAUNCH_CHECK();
                src = BaSrc<CB>{out, nullptr};
                off_in = off_out;
            }
            acc_bases = reinterpret_cast<const typename C::Aff*>(src.pts);
            sorted = nullptr;
            offsets = const_cast<uint32_t*>(off_in);
            acc_total_max = (uint32_t)out_max[ba_rounds - 1];
            uint32_t cap2 = (acc_total_max + 15) / 16;
            if (nthreads > cap2) nthreads = cap2 < 1 ? 1 : cap2;
        }
    }
    acc_kernel<<<(nthreads + ACC_THREADS - 1) / ACC_THREADS, ACC_THREADS, acc_smem, st>>>(
        acc_bases, sorted, offsets, pl.NB, nthreads, buckets, ws.head.as<Xyzz>(), ws.tail.as<Xyzz>());
    G16_LAUNCH_CHECK();
    if (stats && stats->acc_end) G16_CUDA(cudaEventRecord(stats->acc_end, st));
    // the latency-bound tail (stitch, bucket reduction, window sums) may run on a high-priority stream so that it
    // slips in between the CTAs of whatever big kernel another MSM has in flight
    if (tail && tail != st) {
        if (!ws.acc_done) { G16_CUDA(cudaEventCreateWithFlags(&ws.acc_done, cudaEventDisableTiming)); G16_CUDA(cudaEventCreateWithFlags(&ws.tail_done, cudaEventDisableTiming)); }
        G16_CUDA(cudaEventRecord(ws.acc_done, st));
        G16_CUDA(cudaStreamWaitEvent(tail, ws.acc_done, 0));
    }
    cudaStream_t main_st = st;
    if (tail) st = tail;
    uint32_t* heavy = ws.heavy.as<uint32_t>();
    G16_CUDA(cudaMemsetAsync(heavy, 0, 4, st));
    k_fixup<C><<<(pl.NB + 127) / 128, 128, 0, st>>>(offsets, pl.NB, nthreads, buckets, ws.head.as<Xyzz>(), ws.tail.as<Xyzz>(), heavy);
    G16_LAUNCH_CHECK();
    {
        constexpr int HT = IS_G2 ? 128 : 256;
        size_t hs = (size_t)HT * sizeof(Xyzz);
        if (hs > 48 * 1024) G16_CUDA(cudaFuncSetAttribute(k_fixup_heavy<C, HT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hs));
        k_fixup_heavy<C, HT><<<sms * 2, HT, hs, st>>>(offsets, pl.NB, nthreads, buckets, ws.head.as<Xyzz>(), ws.tail.as<Xyzz>(), heavy);
        G16_LAUNCH_CHECK();
    }
    {
        // partial: [U: W x red_rb][T of fold level 0: W x ctas_0][T of level 1: W x ctas_1];  wsum: [U_w][T0_w][T1_w]
        Xyzz* partial = ws.partial.as<Xyzz>();
        Xyzz* wsum = ws.wsum.as<Xyzz>();
        WsumArgs wa = {};
        wa.off[0] = 0; wa.cnt[0] = (uint32_t)pl.red_rb;
        uint32_t poff = (uint32_t)pl.W * pl.red_rb;
        const Xyzz* level_in = buckets;
        Xyzz* level_out = ws.fold.as<Xyzz>();
        for (int k = 0; k < pl.nfold; k++) {
            const uint32_t ctas = msm_fold_ctas(pl, k);
            dim3 fgrid(ctas, pl.W);
            k_bucket_fold<C><<<fgrid, MSM_RED_THREADS, 0, st>>>(level_in, pl.fold_n[k], pl.fold_m[k], level_out, partial + poff);
            G16_LAUNCH_CHECK();
            wa.off[1 + k] = poff; wa.cnt[1 + k] = ctas;
            poff += (uint32_t)pl.W * ctas;
            level_in = level_out;
            level_out += (size_t)pl.W * pl.fold_n[k + 1];
        }
        dim3 rgrid(pl.red_rb, pl.W);
        k_bucket_reduce<C><<<rgrid, MSM_RED_THREADS, 0, st>>>(level_in, pl.fold_n[pl.nfold], pl.red_m, partial);
        G16_LAUNCH_CHECK();
        const size_t wsm = (size_t)MSM_WSUM_THREADS * sizeof(Xyzz);
        if (wsm > 48 * 1024) G16_CUDA(cudaFuncSetAttribute(k_window_sum<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsm));
        dim3 wgrid(pl.W, 1 + pl.nfold);
        k_window_sum<C><<<wgrid, MSM_WSUM_THREADS, wsm, st>>>(partial, wa, wsum);
        G16_LAUNCH_CHECK();
    }
    G16_CUDA(cudaMemcpyAsync(ws.host_wsum.p, ws.wsum.p, (size_t)(1 + pl.nfold) * pl.W * sizeof(Xyzz), cudaMemcpyDeviceToHost, st));
    if (st != main_st) {   // the MSM's own stream completes when its tail does
        G16_CUDA(cudaEventRecord(ws.tail_done, st));
        G16_CUDA(cudaStreamWaitEvent(main_st, ws.tail_done, 0));
    }
    if (stats) stats->acc_threads = nthreads;
}

// Horner over the window sums (host, O(W*c) group operations) -> affine
template <class C>
typename C::Xyzz msm_finish_xyzz(const MsmWorkspace& ws, const MsmPlan& pl) {
    typedef typename C::Xyzz Xyzz;
    const Xyzz* wsum = ws.host_wsum.as<Xyzz>();
    Xyzz acc = C::inf();
    for (int w = pl.W - 1; w >= 0; w--) {
        for (int i = 0; i < pl.c; i++) acc = C::dbl(acc);
        Xyzz u = wsum[w];                                  // set total R_0, R_k = fold_m[k] * R_(k+1) - T_k, R_nfold = U_w
        for (int k = pl.nfold - 1; k >= 0; k--) {
            for (int m = pl.fold_m[k]; m > 1; m >>= 1) u = C::dbl(u);
            C::add(u, C::neg(wsum[(size_t)(1 + k) * pl.W + w]));
        }
        C::add(acc, u);
    }
    return acc;
}
template <class C>
typename C::Aff msm_finish(const MsmWorkspace& ws, const MsmPlan& pl) { return C::to_aff(msm_finish_xyzz<C>(ws, pl)); }

}  // namespace g16
