// Fr NTT kernels and the computeH pipeline for sm_100a.
//
// Replaces, on the reference's accelerated path (backend/groth16/bn254/icicle/icicle.go):
//   :68,73   GenerateTwiddleFactors        -> NttDomain::table (per-stage twiddles, coset shift folded in)
//   :489     INttOnDevice(.., false) x3    -> ntt_dif(TW_INV)           (1/n folded into the pointwise step)
//   :490     NttOnDevice(.., coset) x3     -> ntt_dit(TW_COSET_FWD)     (no CosetTable sweep)
//   :500     PolyOps                       -> fr_pointwise_h            (one fused kernel, scalar den)
//   :502     INttOnDevice(.., coset)       -> ntt_dif(TW_COSET_INV)     (no CosetTableInv sweep)
//   :510     ReverseScalars                -> eliminated: DIF output *is* the bit-reversed order pk.G1.Z uses
// and mirrors the CPU semantics of backend/groth16/bn254/prove.go:356-399 (fft.Domain DIF/DIT).
//
// Kernel shape: one CTA stages a tile of 2^k rows x 2^cb columns (row stride 2^s elements) in shared
// memory as two 128-bit planes, runs k radix-2 stages on it and writes it back; a size-n transform is
// ceil(log n / k) such sweeps.  32-byte elements are read and written as 2 x 128-bit accesses.
#include "ntt.cuh"
#include <cstdlib>

namespace g16 {

// primitive 2^28-th root of unity and coset shift of fft.Domain (gnark-crypto: generator^((r-1)/2^28), generator =
// FrMultiplicativeGen = 5 for BN254, 22 for BLS12-377 -- whose 2^47-th root squared 19 times is the value below)
#if G16_CURVE == 377
static const uint32_t ROOT_2_28[8] = {0xe5f8848au, 0x270a2326u, 0xa727567du, 0x97d14afau, 0x48746fc7u, 0x1a3a5a4eu, 0xa42f077au, 0x0044e4b1u};
static const uint64_t COSET_GEN = 22;
#else
static const uint32_t ROOT_2_28[8] = {0x725b19f0u, 0x9bd61b6eu, 0x41112ed4u, 0x402d111eu, 0x8ef62abcu, 0x00e0a7ebu, 0xa58a7e85u, 0x2a3c09f0u};
static const uint64_t COSET_GEN = 5;
#endif

// tile shape, tuned on B200 (experiments/ntt_sweep.py): 2^10 elements = 32 KB of shared memory per CTA, 4 columns
int ntt_log_tile() { return 10; }
int ntt_log_cols() { return 2; }

// ------------------------------------------------------------------ twiddle tables
struct StageParams {
    Fr::T root[28];   // per stage s: multiplicative step
    Fr::T shift[28];  // per stage s: constant factor
};

__global__ void k_build_table(Fr::T* out, uint64_t n, StageParams sp) {
    // thread t fills E = e+1 in [8t, 8t+8); stage s = floor(log2 E), j = E - 2^s
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t E0 = t * 8;
    if (E0 >= n) return;
    Fr::T val = Fr::zero();
    bool have = false;
    for (int i = 0; i < 8; i++) {
        uint64_t E = E0 + i;
        if (E == 0 || E >= n) continue;
        int s = 63 - __clzll((long long)E);
        uint64_t j = E - (1ull << s);
        if (!have || j == 0) {
            val = Fr::mul(sp.shift[s], Fr::pow_u64(sp.root[s], j));
            have = true;
        } else {
            val = Fr::mul(val, sp.root[s]);
        }
        out[E - 1] = val;
    }
}

void NttDomain::init(int logn_) {
    logn = logn_;
    n = 1ull << logn;
    if (logn > 28) throw std::runtime_error("NTT domain larger than 2^28 (Fr two-adicity)");
    Fr::T root; memcpy(root.l, ROOT_2_28, 32);
    root = Fr::to_mont(root);
    for (int i = logn; i < 28; i++) root = Fr::sqr(root);
    omega = root;
    omega_inv = Fr::inv(root);
    coset_gen = Fr::from_u64(COSET_GEN);
    coset_inv = Fr::inv(coset_gen);
    n_inv = Fr::inv(Fr::from_u64(n));
    Fr::T gn = Fr::pow_u64(coset_gen, n);
    Fr::T den = Fr::inv(Fr::sub(gn, Fr::one()));
    h_c1 = n_inv;
    h_c2 = Fr::mul(den, Fr::sqr(n_inv));
}

// fills p[0, n-1) with table t of domain d (layout in ntt.cuh); enqueues one kernel, allocates nothing
static void fill_table(const NttDomain& d, NttTable t, Fr::T* p, cudaStream_t st) {
    StageParams sp;
    Fr::T w = (t == TW_FWD || t == TW_COSET_FWD) ? d.omega : d.omega_inv;
    Fr::T g = (t == TW_COSET_FWD) ? d.coset_gen : (t == TW_COSET_INV ? d.coset_inv : Fr::one());
    // stage s = logn-1 uses w and g; each lower stage squares both
    for (int s = d.logn - 1; s >= 0; s--) {
        sp.root[s] = w; sp.shift[s] = g;
        w = Fr::sqr(w); g = Fr::sqr(g);
    }
    for (int s = d.logn; s < 28; s++) { sp.root[s] = Fr::one(); sp.shift[s] = Fr::one(); }
    uint64_t threads = (d.n + 7) / 8;
    k_build_table<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(p, d.n, sp);
    G16_LAUNCH_CHECK();
}

const Fr::T* NttDomain::table(NttTable t, cudaStream_t st) {
    if (tw[t]) return tw[t];
    if (logn == 0) { G16_CUDA(cudaMalloc(&tw[t], 32)); return tw[t]; }
    Fr::T* p = nullptr;
    G16_CUDA(cudaMalloc(&p, (size_t)n * sizeof(Fr::T)));
    fill_table(*this, t, p, st);
    tw[t] = p;
    return p;
}

void NttDomain::release() {
    for (int i = 0; i < 4; i++) { if (tw[i]) cudaFree(tw[i]); tw[i] = nullptr; }
}

// ------------------------------------------------------------------ butterfly sweeps
struct PassArgs {
    Fr::T* p[3];
    const Fr::T* tw;
    int s;   // log2 of the row stride (= log2 of the smallest half-size handled by this sweep)
    int k;   // stages in this sweep (tile rows = 2^k)
    int cb;  // log2 of tile columns
    int g;   // distributed sweeps on the D(rb) distribution: log2(world); the rank owns the positions whose bits
    int rb;  //   [rb, rb+g) equal `rank`, local index q = those positions with the rank bits squeezed out
    uint32_t rank;   //   (rb = 0: the cyclic distribution, local q stands for global position q*2^g + rank)
};

__device__ __forceinline__ Fr::T ld_planes(const uint4* lo, const uint4* hi, int e) {
    Fr::T r; uint4 a = lo[e], b = hi[e];
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w; r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
__device__ __forceinline__ void st_planes(uint4* lo, uint4* hi, int e, const Fr::T& v) {
    lo[e] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
    hi[e] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
}
__device__ __forceinline__ Fr::T ldg_fr(const Fr::T* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    Fr::T r;
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w; r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}

constexpr int NTT_THREADS = 256;

template <bool DIF>
__global__ void __launch_bounds__(NTT_THREADS) k_ntt_pass(PassArgs a) {
    extern __shared__ uint4 sm[];
    const int T = 1 << (a.k + a.cb);
    uint4* lo = sm;
    uint4* hi = sm + T;
    Fr::T* data = a.p[blockIdx.y];
    const int s = a.s, k = a.k, cb = a.cb;
    const int cmask = (1 << cb) - 1;
    const uint64_t qbase = (uint64_t)blockIdx.x << cb;
    const uint64_t smask = (1ull << s) - 1;

    for (int e = threadIdx.x; e < T; e += NTT_THREADS) {
        int c = e & cmask, r = e >> cb;
        uint64_t q = qbase + c;
        uint64_t gi = ((q >> s) << (s + k)) | ((uint64_t)r << s) | (q & smask);
        const uint4* src = reinterpret_cast<const uint4*>(data + gi);
        lo[e] = src[0];
        hi[e] = src[1];
    }
    __syncthreads();

    const int nb = T >> 1;
    for (int st = 0; st < k; st++) {
        const int m = DIF ? (k - 1 - st) : st;
        const int mmask = (1 << m) - 1;
        const Fr::T* tws = a.tw + ((1ull << (m + s + a.g)) - 1);
        for (int bi = threadIdx.x; bi < nb; bi += NTT_THREADS) {
            int c = bi & cmask, rb = bi >> cb;
            int r0 = ((rb >> m) << (m + 1)) | (rb & mmask);
            int e0 = (r0 << cb) | c;
            int e1 = e0 + (1 << (m + cb));
            uint64_t q = qbase + c;
            uint64_t j = ((uint64_t)(r0 & mmask) << s) | (q & smask);
            const uint64_t jg = ((j >> a.rb) << (a.rb + a.g)) | ((uint64_t)a.rank << a.rb) | (j & ((1ull << a.rb) - 1));
            Fr::T t = ldg_fr(tws + jg);
            // butterflies in the lazily reduced field (values in [0, 2r) while the tile is in shared memory)
            Fr::T u = ld_planes(lo, hi, e0), v = ld_planes(lo, hi, e1);
            if (DIF) {
                st_planes(lo, hi, e0, FrL::add(u, v));
                st_planes(lo, hi, e1, FrL::mul(FrL::sub(u, v), t));
            } else {
                v = FrL::mul(v, t);
                st_planes(lo, hi, e0, FrL::add(u, v));
                st_planes(lo, hi, e1, FrL::sub(u, v));
            }
        }
        __syncthreads();
    }

    for (int e = threadIdx.x; e < T; e += NTT_THREADS) {
        int c = e & cmask, r = e >> cb;
        uint64_t q = qbase + c;
        uint64_t gi = ((q >> s) << (s + k)) | ((uint64_t)r << s) | (q & smask);
        uint4* dst = reinterpret_cast<uint4*>(data + gi);
        Fr::T v = FrL::canon(ld_planes(lo, hi, e));     // back to [0, r): global memory only ever holds canonical values
        dst[0] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
        dst[1] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
    }
}

struct Sweep { int s, k, cb; };

// sweeps covering the stages whose (local) log2 half-size lies in [lo, hi), bottom-up
static std::vector<Sweep> plan_range(int lo, int hi) {
    std::vector<Sweep> v;
    const int nst = hi - lo;
    if (nst <= 0) return v;
    const int LT = ntt_log_tile();
    if (lo == 0 && nst <= LT) { v.push_back({0, nst, 0}); return v; }
    const int cb = ntt_log_cols();
    const int kmax = LT - cb;
    const int total = (nst + kmax - 1) / kmax;
    std::vector<int> ks(total, nst / total);
    for (int i = 0; i < nst % total; i++) ks[i]++;
    int s = lo;
    for (int i = 0; i < total; i++) {
        int c = cb < s ? cb : s;                // keep a tile row contiguous in memory
        while (ks[i] + c > LT) c--;
        v.push_back({s, ks[i], c});
        s += ks[i];
    }
    return v;
}
static std::vector<Sweep> plan_sweeps(int logn) { return plan_range(0, logn); }

template <bool DIF>
static void launch_sweep(uint64_t n_local, const NttBatch& b, const Fr::T* tw, const Sweep& sw, cudaStream_t st, int g = 0, uint32_t rank = 0, int rb = 0) {
    PassArgs a;
    for (int i = 0; i < 3; i++) a.p[i] = b.p[i < b.count ? i : 0];
    a.tw = tw; a.s = sw.s; a.k = sw.k; a.cb = sw.cb; a.g = g; a.rank = rank; a.rb = rb;
    size_t smem = (size_t)2 * sizeof(uint4) << (sw.k + sw.cb);
    if (smem > 48 * 1024)
        G16_CUDA(cudaFuncSetAttribute(k_ntt_pass<DIF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    uint64_t gx = (n_local >> sw.k) >> sw.cb;
    dim3 grid((unsigned)gx, (unsigned)b.count);
    k_ntt_pass<DIF><<<grid, NTT_THREADS, smem, st>>>(a);
    G16_LAUNCH_CHECK();
}

void ntt_dif(NttDomain& d, const NttBatch& b, NttTable t, cudaStream_t st) {
    if (d.logn == 0) return;
    const Fr::T* tw = d.table(t, st);
    std::vector<Sweep> v = plan_sweeps(d.logn);
    for (int i = (int)v.size() - 1; i >= 0; i--) launch_sweep<true>(d.n, b, tw, v[i], st);
}
void ntt_dit(NttDomain& d, const NttBatch& b, NttTable t, cudaStream_t st) {
    if (d.logn == 0) return;
    const Fr::T* tw = d.table(t, st);
    std::vector<Sweep> v = plan_sweeps(d.logn);
    for (size_t i = 0; i < v.size(); i++) launch_sweep<false>(d.n, b, tw, v[i], st);
}

// ------------------------------------------------------------------ element-wise kernels
__global__ void k_scale_powers(Fr::T* a, uint64_t n, int logn, Fr::T c, Fr::T base, int bitrev_index, int use_base) {
    // thread handles 16 consecutive exponents
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t i0 = t * 16;
    if (i0 >= n) return;
    Fr::T p = c;
    if (use_base) p = Fr::mul(c, Fr::pow_u64(base, i0));
    for (int k = 0; k < 16 && i0 + k < n; k++) {
        uint64_t i = i0 + k;
        uint64_t pos = bitrev_index ? (__brevll(i) >> (64 - logn)) : i;
        if (logn == 0) pos = 0;
        a[pos] = Fr::mul(a[pos], p);
        if (use_base) p = Fr::mul(p, base);
    }
}
void fr_scale_powers(Fr::T* a, uint64_t n, int logn, const Fr::T& c, const Fr::T& base, int bitrev_index, int use_base, cudaStream_t st) {
    uint64_t threads = (n + 15) / 16;
    k_scale_powers<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(a, n, logn, c, base, bitrev_index, use_base);
    G16_LAUNCH_CHECK();
}

__global__ void k_pointwise_h(Fr::T* a, const Fr::T* b, const Fr::T* c, uint64_t n, Fr::T c1, Fr::T c2) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        Fr::T x = ldg_fr(a + i), y = ldg_fr(b + i), z = ldg_fr(c + i);
        Fr::T t = Fr::mul(Fr::mul(x, y), c1);
        a[i] = Fr::mul(Fr::sub(t, z), c2);
    }
}
void fr_pointwise_h(Fr::T* a, const Fr::T* b, const Fr::T* c, uint64_t n, const Fr::T& c1, const Fr::T& c2, cudaStream_t st) {
    unsigned blocks = (unsigned)((n + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_pointwise_h<<<blocks, 256, 0, st>>>(a, b, c, n, c1, c2);
    G16_LAUNCH_CHECK();
}

__global__ void k_mul_vec(Fr::T* out, const Fr::T* a, const Fr::T* b, uint64_t n) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        out[i] = Fr::mul(ldg_fr(a + i), ldg_fr(b + i));
}
void fr_mul_vec(Fr::T* out, const Fr::T* a, const Fr::T* b, uint64_t n, cudaStream_t st) {
    if (!n) return;
    unsigned blocks = (unsigned)((n + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_mul_vec<<<blocks, 256, 0, st>>>(out, a, b, n);
    G16_LAUNCH_CHECK();
}

// ------------------------------------------------------------------ distributed transforms (SURVEY.md 8(e) level 3)
// A size-n transform over G = 2^g ranks is split at a global stage X: stages with half-size >= 2^X run on the
// CYCLIC distribution (rank r holds positions q*G + r; pairs stay local because the half-size is a multiple of G),
// stages below X run on the BLOCK distribution (rank r holds [r n/G, (r+1) n/G)).  One all-to-all between the two
// (the four-step transpose) -- the twiddles need no extra pass because every stage has its own table.
// rb generalises "cyclic" (rb = 0) to the distribution D(rb): the rank is given by the position bits [rb, rb+g).  Every stage
// with half-size 2^t, t outside [rb, rb+g), is local under D(rb); the block distribution is D(L-g).  With rb = split - g the
// upper phase runs on runs of 2^rb consecutive positions, so host arrays can be loaded straight into it (strided copies of
// 2^rb x 32 B) and the exchanges move contiguous runs (multi.cu).
void ntt_dist_stages(NttDomain& d, const NttBatch& b, NttTable t, bool dif, int world_log, uint32_t rank, bool cyclic, int split, cudaStream_t st, int rb) {
    const Fr::T* tw = d.table(t, st);
    const int L = d.logn, g = world_log, Ll = L - g;
    if (split < g || split > Ll) throw std::runtime_error("distributed NTT: split stage must lie in [log2(world), log2(n/world)]");
    if (rb < 0 || rb + g > split) throw std::runtime_error("distributed NTT: rank bits must lie below the split stage");
    const uint64_t n_local = d.n >> g;
    std::vector<Sweep> v = cyclic ? plan_range(split - g, Ll) : plan_range(0, split);
    if (dif) for (int i = (int)v.size() - 1; i >= 0; i--) launch_sweep<true>(n_local, b, tw, v[i], st, cyclic ? g : 0, cyclic ? rank : 0, cyclic ? rb : 0);
    else for (size_t i = 0; i < v.size(); i++) launch_sweep<false>(n_local, b, tw, v[i], st, cyclic ? g : 0, cyclic ? rank : 0, cyclic ? rb : 0);
}

// The exchange between the two phases of a distributed transform, fused with both layout changes: every element of the
// local arrays is stored straight into its final place on the GPU that owns it in the other distribution (peer memory
// over NVLink / NVSwitch; the local share is an ordinary store).  Runs of 2^rb elements stay contiguous on both sides.
__global__ void __launch_bounds__(256) k_dist_exchange(DistExchange a) {
    const uint64_t total = (uint64_t)a.count * a.n_local;
    const uint64_t rmask = (1ull << a.rb) - 1, lmask = a.n_local - 1;
    const int Ll = a.L - a.g;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t arr = e / a.n_local, q = e & lmask;
        uint64_t dq; uint32_t dr;
        if (a.to_block) {   // D(rb) -> block
            const uint64_t i = ((q >> a.rb) << (a.rb + a.g)) | ((uint64_t)a.rank << a.rb) | (q & rmask);
            dr = (uint32_t)(i >> Ll); dq = i & lmask;
        } else {            // block -> D(rb)
            const uint64_t i = ((uint64_t)a.rank << Ll) | q;
            dr = (uint32_t)(i >> a.rb) & ((1u << a.g) - 1); dq = ((i >> (a.rb + a.g)) << a.rb) | (i & rmask);
        }
        const uint4* src = reinterpret_cast<const uint4*>(a.src + e);
        uint4* dst = reinterpret_cast<uint4*>(a.dst[dr] + arr * a.n_local + dq);
        const uint4 lo = src[0], hi = src[1];
        dst[0] = lo; dst[1] = hi;
    }
}
void ntt_dist_exchange(const DistExchange& a, cudaStream_t st) {
    const uint64_t total = (uint64_t)a.count * a.n_local;
    unsigned blocks = (unsigned)((total + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_dist_exchange<<<blocks, 256, 0, st>>>(a);
    G16_LAUNCH_CHECK();
}

// wire layout of an all-to-all: [peer x][array][q'], q' < m = n_local / G.
//   cyclic-local index of (x, q') = x*m + q'      block-local index of (x, q') = q'*G + x
__global__ void k_dist_permute(const uint4* __restrict__ src, uint4* __restrict__ dst, int count, uint64_t n_local, uint32_t G,
                               int block_layout, int to_wire) {
    const uint64_t m = n_local / G, total = (uint64_t)count * n_local;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t qp = e % m, rest = e / m;
        uint32_t arr = (uint32_t)(rest % count), x = (uint32_t)(rest / count);
        uint64_t local = (uint64_t)arr * n_local + (block_layout ? qp * G + x : (uint64_t)x * m + qp);
        uint64_t from = to_wire ? local : e, to = to_wire ? e : local;
        dst[2 * to] = src[2 * from];
        dst[2 * to + 1] = src[2 * from + 1];
    }
}
void ntt_dist_permute(const Fr::T* src, Fr::T* dst, int count, uint64_t n_local, uint32_t world, int block_layout, int to_wire, cudaStream_t st) {
    uint64_t total = (uint64_t)count * n_local;
    unsigned blocks = (unsigned)((total + 255) / 256);
    if (blocks > 148 * 32) blocks = 148 * 32;
    k_dist_permute<<<blocks, 256, 0, st>>>(reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), count, n_local, world, block_layout, to_wire);
    G16_LAUNCH_CHECK();
}

// ------------------------------------------------------------------ computeH (prove.go:356-399 / icicle.go:453-513)
void compute_h_device(NttDomain& d, Fr::T* a, Fr::T* b, Fr::T* c, cudaStream_t st) {
    NttBatch abc{{a, b, c}, 3};
    ntt_dif(d, abc, TW_INV, st);          // n * coefficients, bit-reversed
    ntt_dit(d, abc, TW_COSET_FWD, st);    // n * evaluations on g<w>, natural order
    fr_pointwise_h(a, b, c, d.n, d.h_c1, d.h_c2, st);
    NttBatch one{{a, a, a}, 1};
    ntt_dif(d, one, TW_COSET_INV, st);    // h coefficients, bit-reversed order
}

// ------------------------------------------------------------------ computeH without any device allocation
// For callers that may not allocate (CUDA graph capture, tight serving loops): the padded working copies of a, b, c
// and the three twiddle tables live in a file-scope device array instead of cudaMalloc'ed buffers.  The array is
// shared by every call on the device (calls that overlap on the GPU race) and keeps the previous call's contents:
// everything a call reads from it is (re)written by that call first -- the tables are rebuilt, the padding rows
// [n_constraints, n) are cleared.  No counterpart in the reference.
constexpr uint64_t NOALLOC_N = 1ull << 16;                 // G16_NOALLOC_MAX_DOMAIN (include/g16b200.h)
// [a | b | c | TW_INV | TW_COSET_FWD | TW_COSET_INV], NOALLOC_N Fr each: 12 MiB per device (per curve)
static __device__ __align__(256) unsigned char g_noalloc_scratch[6 * NOALLOC_N * sizeof(Fr::T)];

// device address of the array, resolved once per device with that device current (in host code the array's name is
// only a stand-in for the symbol, not an address a kernel can use)
static Fr::T* noalloc_scratch(int device) {
    static std::mutex mu;
    static Fr::T* ptr[64] = {};
    if (device < 0 || device >= 64) throw std::runtime_error("compute_h_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[device]) {
        void* p = nullptr;
        G16_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        ptr[device] = reinterpret_cast<Fr::T*>(p);
    }
    return ptr[device];
}

void compute_h_noalloc(int device, const Fr::T* a, const Fr::T* b, const Fr::T* c, uint64_t n_constraints, uint64_t n,
                       Fr::T* h_out, cudaStream_t st) {
    if (!n || (n & (n - 1))) throw std::runtime_error("compute_h_noalloc: domain_n must be a power of two");
    if (n > NOALLOC_N) throw std::runtime_error("compute_h_noalloc: domain_n " + std::to_string(n) + " exceeds the static scratch (maximum 2^16 = " + std::to_string(NOALLOC_N) + "); use g16_compute_h_dev");
    if (n_constraints > n) throw std::runtime_error("compute_h_noalloc: more constraints than the domain holds");
    Fr::T* base = noalloc_scratch(device);
    NttDomain d;                     // host constants only; its tables point into the static array
    d.init(ilog2_ceil(n));
    const Fr::T* src[3] = {a, b, c};
    Fr::T* w[3] = {base, base + NOALLOC_N, base + 2 * NOALLOC_N};
    for (int i = 0; i < 3; i++) {
        if (n_constraints) G16_CUDA(cudaMemcpyAsync(w[i], src[i], (size_t)n_constraints * sizeof(Fr::T), cudaMemcpyDeviceToDevice, st));
        if (n > n_constraints) G16_CUDA(cudaMemsetAsync(w[i] + n_constraints, 0, (size_t)(n - n_constraints) * sizeof(Fr::T), st));
    }
    const NttTable tabs[3] = {TW_INV, TW_COSET_FWD, TW_COSET_INV};
    for (int i = 0; i < 3; i++) {
        d.tw[tabs[i]] = base + (3 + i) * NOALLOC_N;
        if (d.logn > 0) fill_table(d, tabs[i], d.tw[tabs[i]], st);
    }
    compute_h_device(d, w[0], w[1], w[2], st);
    G16_CUDA(cudaMemcpyAsync(h_out, w[0], (size_t)n * sizeof(Fr::T), cudaMemcpyDeviceToDevice, st));
}

}  // namespace g16
