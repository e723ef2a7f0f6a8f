// The context object behind the C ABI (include/g16b200.h) and what api.cu and multi.cu share.
//
// A g16_ctx drives one GPU (streams, workspaces, resident keys) -- or, created over several devices, a group of
// such single-GPU contexts ("ranks", multi.cu) that one g16_prove_bn254 call uses together: the reference surface is
// ONE in-process call (backend/groth16/bn254/icicle/icicle.go:133), so the multi-GPU choreography lives behind it.
#pragma once
#include "curve_select.h"
#include "../../include/g16b200.h"
#include "msm_api.cuh"
#include "ntt.cuh"
#include "r1cs.cuh"
#include "team.h"

#include <cstring>
#include <memory>

namespace g16 {

static constexpr size_t G1SZ = sizeof(G1::Aff), G2SZ = sizeof(G2::Aff);   // 64 / 128 B (BN254), 96 / 192 B (BLS12-377)

// the slice of the key one context keeps: element range [lo, hi) of each base array, in G16_OUT_* order
// (AR = G1.A, BS1 = G1.B, KRS_K = G1.K, KRS_HZ = G1.Z, BS2 = G2.B)
struct PkRanges { uint64_t lo[5], hi[5]; };

struct PkDev {
    uint64_t domain_n = 0;
    int logn = 0;
    uint64_t n_wires = 0;
    // this shard's slices
    DevBuf g1_a, g1_b, g1_k, g1_z, g2_b;
    DevBuf map_a, map_b, map_k, map_b2;   // u32 wire index per base (map_b2 only when the G1.B and G2.B slices differ)
    uint64_t n_a = 0, n_b = 0, n_k = 0, n_z = 0, n_b2 = 0;
    uint64_t k_pad = 0;          // points at infinity in front of G1.K (included in n_k) that line it up with G1.A
    bool same_ak = false;        // ... so that the K MSM reuses the A MSM's sort
    int pre_c[5] = {0, 0, 0, 0, 0}; // window bits of the precomputed-multiples table per MSM (G16_OUT_* order); 0 = none
    PkRanges ranges = {};         // the slices as requested at upload
    uint64_t z_off = 0;           // first h index of this shard's Z slice
    uint64_t w_lo = 0, w_hi = 0;  // wires [w_lo, w_hi) are the only ones this shard's A / B / K maps read
    bool same_b = true;           // G1.B and G2.B slices cover the same range: the bs1 MSM reuses the bs2 sort
    void release() {
        g1_a.release(); g1_b.release(); g1_k.release(); g1_z.release(); g2_b.release();
        map_a.release(); map_b.release(); map_k.release(); map_b2.release();
    }
};

struct MultiGpu;   // multi.cu

// Caller memory that is not page-locked (a Go caller's heap slices: runtime.Pinner pins against the garbage collector,
// it does not page-lock) cannot be the source of an asynchronous DMA: cudaMemcpyAsync stages it through the driver's
// own bounce buffer on the calling thread, measured ~10 GB/s.  The stager does that copy itself with several threads
// into a ring of pinned buffers and issues the DMA of one chunk while the next is being gathered.
struct HostStager {
    static constexpr size_t CHUNK = (size_t)8 << 20;
    static constexpr int NBUF = 4;
    PinnedBuf buf[NBUF];
    cudaEvent_t done[NBUF] = {};
    bool used[NBUF] = {};
    int next = 0;
    Team team;
    int threads = 0;            // 0 = not started

    static bool pageable(const void* p) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
        return a.type == cudaMemoryTypeUnregistered;
    }
    void start(int want_threads) {
        if (threads) return;
        threads = want_threads < 1 ? 1 : (want_threads > 16 ? 16 : want_threads);
        for (int i = 0; i < NBUF; i++) { buf[i].reserve(CHUNK); G16_CUDA(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming)); }
        team.start(threads);
    }
    void release() {
        if (threads) team.shutdown();
        threads = 0;
        for (int i = 0; i < NBUF; i++) { buf[i].release(); if (done[i]) cudaEventDestroy(done[i]); done[i] = nullptr; used[i] = false; }
    }
    // `rows` rows of row_bytes each, source pitch src_pitch (contiguous: rows = 1), packed densely at dst_d
    void copy_rows(void* dst_d, const void* src_h, size_t row_bytes, size_t src_pitch, size_t rows, cudaStream_t st) {
        if (!rows || !row_bytes) return;
        const char* src = (const char*)src_h;
        char* dst = (char*)dst_d;
        if (rows == 1 || src_pitch == row_bytes) { src_pitch = row_bytes = row_bytes * rows; rows = 1; }
        // pieces of at most CHUNK bytes: whole rows when they fit, else slices of one row
        const size_t rows_per = row_bytes >= CHUNK ? 1 : CHUNK / row_bytes;
        for (size_t r0 = 0; r0 < rows; r0 += rows_per) {
            const size_t nr = rows - r0 < rows_per ? rows - r0 : rows_per;
            for (size_t off = 0; off < row_bytes; off += CHUNK) {
                const size_t len = row_bytes - off < CHUNK ? row_bytes - off : CHUNK;      // only < row_bytes when nr == 1
                const size_t bytes = nr == 1 ? len : nr * row_bytes;
                const int b = next++ % NBUF;
                if (used[b]) G16_CUDA(cudaEventSynchronize(done[b]));
                char* stage = buf[b].as<char>();
                const int T = threads;
                team.run([&](int t) {
                    if (nr == 1) {
                        const size_t lo = len * t / T, hi = len * (t + 1) / T;
                        memcpy(stage + lo, src + r0 * src_pitch + off + lo, hi - lo);
                    } else {
                        for (size_t r = nr * t / T; r < nr * (t + 1) / T; r++) memcpy(stage + r * row_bytes, src + (r0 + r) * src_pitch, row_bytes);
                    }
                });
                G16_CUDA(cudaMemcpyAsync(dst + r0 * row_bytes + off, stage, bytes, cudaMemcpyHostToDevice, st));
                G16_CUDA(cudaEventRecord(done[b], st));
                used[b] = true;
                if (nr > 1) break;
            }
        }
    }
    void copy(void* dst_d, const void* src_h, size_t bytes, cudaStream_t st) { copy_rows(dst_d, src_h, bytes, bytes, 1, st); }
};

}  // namespace g16

struct g16_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t aux[4] = {};   // the MSMs that do not depend on h
    cudaStream_t tails[5] = {}; // high-priority streams for the latency-bound MSM tails
    int streams = 5;
    cudaEvent_t ev[32] = {};
    std::string err;
    std::map<int, g16::NttDomain> domains;
    std::map<uint64_t, g16::PkDev> pks;
    std::map<uint64_t, g16::R1csDev> r1cs;
    struct Bases { g16::DevBuf table; uint64_t n = 0; int pre_c = 0; bool g2 = false; };
    std::map<uint64_t, Bases> bases;
    struct CommitKey { Bases basis, basis_exp_sigma; };   // pedersen.ProvingKey (setup.go:57)
    std::map<uint64_t, CommitKey> commit_keys;
    uint64_t next_pk = 1;
    g16::MsmPlan pend_plan[5];   // g16_prove_msms_begin_dev .. g16_prove_msms_dev: plans of the MSMs already in flight
    uint64_t pend_pk = 0;        // 0 = nothing pending
    g16::MsmWorkspace ws[5];
    g16::DevBuf stage[4];        // host-pointer variants: staged inputs
    g16::PinnedBuf file_stage[2];// g16_pk_upload_from_file: double-buffered pread -> H2D
    g16::DevBuf fb_table_g1, fb_table_g2;
    int msm_c = 0, msm_waves = 4;
    int dist_sync = 1;           // g16_dist_*_dev return after the stream has drained (0: stream-ordered, caller synchronises)
    int msm_precompute = 1;      // keep 2^(c j) multiples of every key base resident (one shared bucket set, larger windows)
    int msm_precompute_g1 = 1;   // ... for the G1 arrays too (0: only G2.B gets a table; for keys that fill the HBM)
    int msm_pre_c = 0;           // window bits of the tables built by the next uploads (0 = chosen from the array size)
    int g2_weight = 0;           // multi-GPU placement: cost of a G2 point in hundredths of a G1 point (0 = by group size)
    int stage_threads = 0;       // copier threads for pageable caller memory (0 = from the core count)
    g16::HostStager stager;
    g16::MultiGpu* multi = nullptr;   // non-null on a context created over several devices (this context is rank 0)

    g16::NttDomain& domain(int logn) {
        auto it = domains.find(logn);
        if (it != domains.end()) return it->second;
        g16::NttDomain d;
        d.init(logn);            // throws for logn > 28: nothing half-initialised is left behind
        return domains.emplace(logn, d).first->second;
    }
};

namespace g16 {

template <class F>
int guarded(g16_ctx* ctx, F&& f) {
    if (!ctx) return 1;
    try {
        G16_CUDA(cudaSetDevice(ctx->device));
        f();
        return 0;
    } catch (const std::exception& e) {
        ctx->err = e.what();
        cudaGetLastError();   // clear a sticky launch error so that later calls report their own
        return 2;
    }
}

inline bool is_pow2(uint64_t n) { return n && !(n & (n - 1)); }

inline void slice(uint64_t len, uint32_t idx, uint32_t cnt, uint64_t& lo, uint64_t& hi) {
    if (cnt <= 1) { lo = 0; hi = len; return; }
    lo = len * idx / cnt;
    hi = len * (idx + 1) / cnt;
}

// host -> device copy of caller memory on `st`: asynchronous DMA when the source is page-locked, staged through the
// context's pinned ring by its copier threads when it is pageable (returns when the last chunk has been handed to the DMA)
inline void h2d_caller(g16_ctx* ctx, void* dst_d, const void* src_h, size_t bytes, cudaStream_t st, int ranks_sharing_host = 1) {
    if (!bytes) return;
    if (bytes < ((size_t)1 << 20) || !HostStager::pageable(src_h)) { G16_CUDA(cudaMemcpyAsync(dst_d, src_h, bytes, cudaMemcpyHostToDevice, st)); return; }
    if (!ctx->stager.threads) {
        int t = ctx->stage_threads;
        if (t <= 0) { t = (int)std::thread::hardware_concurrency() / (2 * ranks_sharing_host); t = t < 2 ? 2 : (t > 8 ? 8 : t); }
        ctx->stager.start(t);
    }
    ctx->stager.copy(dst_d, src_h, bytes, st);
}

inline void upload(DevBuf& dst, const void* src, size_t bytes, cudaStream_t st) {
    dst.reserve(bytes ? bytes : 16);
    if (bytes) G16_CUDA(cudaMemcpyAsync(dst.p, src, bytes, cudaMemcpyHostToDevice, st));
}

template <class Aff>
void copy_out(void* dst, const Aff& p) { memcpy(dst, &p, sizeof(Aff)); }

// Where key material comes from: host memory, or a byte range of a file (a ProvingKey.WriteDump image, whose point
// slices are raw memory = the device format; marshal.go:386-456).
struct Src {
    const char* host = nullptr;
    int fd = -1;
    uint64_t off = 0;
    Src at(uint64_t bytes) const { Src s = *this; if (host) s.host += bytes; s.off += bytes; return s; }
};

// ---- api.cu
// creates / destroys a single-GPU context (no group)
g16_ctx* ctx_create_single(int device);
void ctx_destroy_single(g16_ctx* ctx);
// uploads the slices `rg` of the key described by d (src: G1.A, G1.B, G1.K, G1.Z, G2.B) under `handle`; throws
void pk_upload_ranges(g16_ctx* ctx, const g16_pk_desc* d, const Src* src, const PkRanges& rg, uint64_t handle);
void pk_validate(const g16_pk_desc* d);
PkRanges pk_shard_ranges(const g16_pk_desc* d);
// The device section of a prove on one context.  phase 0: everything.  phase 1: enqueue the four MSMs over W and return
// (they run while the caller produces h); phase 2: enqueue the H.Z MSM, wait for all five, finish.
// partial: when non-null receives the five results as XYZZ (no inversion; multi.cu sums the ranks first) and out_points
// is not written.  acc_gate (phase 1): the bucket accumulations wait for this event, the sorts do not.
struct ProvePartial { G1::Xyzz g1[4]; G2::Xyzz g2; };
void prove_device(g16_ctx* ctx, PkDev& pk, const Fr::T* W_d, Fr::T* a_d, Fr::T* b_d, Fr::T* c_d, void* out_points,
                  g16_stats* stats, uint64_t launches0, cudaEvent_t w_ready, bool h_given = false, int phase = 0,
                  ProvePartial* partial = nullptr, cudaEvent_t acc_gate = nullptr);

// ---- multi.cu
MultiGpu* multi_create(g16_ctx* rank0, const int* device_ids, int n_devices);
void multi_destroy(MultiGpu* mg);
void multi_set(MultiGpu* mg, const char* key, int64_t value);
void multi_pk_upload(g16_ctx* ctx, const g16_pk_desc* d, const Src* src, uint64_t* pk_handle);
void multi_pk_free(g16_ctx* ctx, uint64_t pk_handle);
void multi_pk_wire_range(g16_ctx* ctx, uint64_t pk_handle, uint64_t* lo, uint64_t* hi);
const PkDev& multi_pk_of_rank(g16_ctx* ctx, uint64_t pk_handle, int rank);
void multi_r1cs_upload(g16_ctx* ctx, const g16_r1cs_desc* d, uint64_t handle);
void multi_r1cs_free(g16_ctx* ctx, uint64_t handle);
// inputs of a group prove: host arrays, device arrays on rank 0's GPU, or the wire vector + resident matrices
struct MultiIn {
    const void* W = nullptr; const void* a = nullptr; const void* b = nullptr; const void* c = nullptr;
    uint64_t n_wires = 0, n_constraints = 0;
    bool device = false;        // W, a, b, c are device pointers on rank 0's GPU (a, b, c: domain_n elements)
    uint64_t r1cs = 0;          // non-zero: a, b, c are evaluated on the device from this resident constraint system
};
void multi_prove(g16_ctx* ctx, uint64_t pk_handle, const MultiIn& in, void* out_points, g16_stats* stats);
// group primitives (configs 2 and 3 on N GPUs); arrays named _d live on rank 0's GPU
bool multi_can_distribute(g16_ctx* ctx, uint64_t n);
void multi_compute_h_dev(g16_ctx* ctx, void* a_d, void* b_d, void* c_d, uint64_t n);
bool multi_ntt_dev(g16_ctx* ctx, void* data_d, uint64_t n, int inverse, int coset, int decimation);
void multi_bases_upload(g16_ctx* ctx, const void* bases, uint64_t n, int g2, int from_device, uint64_t* handle);
bool multi_has_bases(g16_ctx* ctx, uint64_t handle);
void multi_bases_free(g16_ctx* ctx, uint64_t handle);
void multi_msm_resident(g16_ctx* ctx, uint64_t handle, const void* scalars, bool on_device, void* out);
// api.cu: bases (n points, host memory or -- from_device -- memory of GPU src_device) resident on ctx under `handle`
void bases_upload_one(g16_ctx* ctx, const void* bases, uint64_t n, bool g2, bool from_device, int src_device, uint64_t handle);

}  // namespace g16
