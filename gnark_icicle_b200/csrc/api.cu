// C ABI of libg16b200 (declared in include/g16b200.h): context, device-resident proving key, the prove
// hot path, primitives and measurement helpers.  Host logic only -- kernels live in ntt.cu / msm*.cu.
//
// Reference shape being replaced: backend/groth16/bn254/icicle/icicle.go (setupDevicePointers :31-130,
// Prove :133-422, computeH :453-513) and icicle/provingkey.go:10-28 (deviceInfo).
#include <chrono>
#include <cstdlib>
#include <memory>
#include <unistd.h>

#include "ctx.cuh"

using namespace g16;

namespace {

// Key bases: row 0 of a table that, when it fits comfortably (less than half of the free HBM) and the array is
// large enough to pay for it, also holds the 2^(c j) multiples used by the shared-bucket MSM (msm.cuh, precomputed
// mode).  pre_c = window bits of the table, 0 = bases only.
constexpr uint64_t PRECOMPUTE_MIN_POINTS = 1 << 14;
Src host_src(const void* p) { Src s; s.host = (const char*)p; return s; }

// file -> HBM without an intermediate pageable copy: pread into two pinned staging buffers, the H2D copy of one
// chunk in flight while the next is being read
constexpr size_t FILE_CHUNK = (size_t)32 << 20;
void copy_in(g16_ctx* ctx, void* dst, const Src& src, size_t bytes, cudaStream_t st) {
    if (!bytes) return;
    if (src.fd < 0) { G16_CUDA(cudaMemcpyAsync(dst, src.host, bytes, cudaMemcpyHostToDevice, st)); return; }
    cudaEvent_t* done = &ctx->ev[28];   // ev[28], ev[29]: staging buffer free again
    for (int b = 0; b < 2; b++) ctx->file_stage[b].reserve(FILE_CHUNK);
    size_t pos = 0;
    for (int it = 0; pos < bytes; it++) {
        const int b = it & 1;
        const size_t len = bytes - pos < FILE_CHUNK ? bytes - pos : FILE_CHUNK;
        if (it >= 2) G16_CUDA(cudaEventSynchronize(done[b]));
        char* buf = ctx->file_stage[b].as<char>();
        size_t got = 0;
        while (got < len) {
            ssize_t r = pread(src.fd, buf + got, len - got, (off_t)(src.off + pos + got));
            if (r < 0) throw std::runtime_error("pread failed while loading the key dump");
            if (r == 0) throw std::runtime_error("key dump is shorter than its slice headers say");
            got += (size_t)r;
        }
        G16_CUDA(cudaMemcpyAsync((char*)dst + pos, buf, len, cudaMemcpyHostToDevice, st));
        G16_CUDA(cudaEventRecord(done[b], st));
        pos += len;
    }
    G16_CUDA(cudaStreamSynchronize(st));   // the staging buffers are reused by the next array
}

// pad_front: that many points at infinity in front of the array (the table then has n + pad_front columns)
template <class C>
void upload_bases(g16_ctx* ctx, DevBuf& dst, const Src& src, uint64_t n, int& pre_c, cudaStream_t st, uint64_t pad_front = 0) {
    constexpr bool IS_G2 = sizeof(typename C::E) > sizeof(Fp::T);
    const uint64_t cols = n + pad_front;
    const size_t row = (size_t)cols * sizeof(typename C::Aff);
    pre_c = 0;
    if (ctx->msm_precompute && (IS_G2 || ctx->msm_precompute_g1) && cols >= PRECOMPUTE_MIN_POINTS) {
        const int c = msm_pre_choose_c(cols, IS_G2, ctx->msm_pre_c);
        const size_t total = row * (size_t)(255 / c + 1);
        size_t free_b = 0, total_b = 0;
        G16_CUDA(cudaMemGetInfo(&free_b, &total_b));
        if (total <= free_b / 2) pre_c = c;
    }
    dst.reserve(pre_c ? row * (size_t)(255 / pre_c + 1) : (row ? row : 16));
    if (pad_front) G16_CUDA(cudaMemsetAsync(dst.p, 0, (size_t)pad_front * sizeof(typename C::Aff), st));   // (0, 0) = infinity
    copy_in(ctx, (char*)dst.p + (size_t)pad_front * sizeof(typename C::Aff), src, (size_t)n * sizeof(typename C::Aff), st);
    if (pre_c) {
        if constexpr (IS_G2) msm_g2_precompute(dst.as<G2::Aff>(), cols, pre_c, st);
        else msm_g1_precompute(dst.as<G1::Aff>(), cols, pre_c, st);
    }
}
template <class C>
void upload_bases(g16_ctx* ctx, DevBuf& dst, const void* src, uint64_t n, int& pre_c, cudaStream_t st) {
    upload_bases<C>(ctx, dst, host_src(src), n, pre_c, st);
}

}  // namespace

namespace g16 {

g16_ctx* ctx_create_single(int device) {
    std::unique_ptr<g16_ctx> c(new g16_ctx());
    // A prove uses 10 streams (main, 4 MSM streams, 5 tail streams); with the default 8 hardware queues two of them
    // share one and the later MSM starts only when the earlier one has finished.  Takes effect only if CUDA has not
    // been initialised in this process yet (hosts that initialise CUDA first should export it themselves).
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
    int count = 0;
    G16_CUDA(cudaGetDeviceCount(&count));
    if (count < 1) throw std::runtime_error("no CUDA device: libg16b200 has no CPU fallback");
    if (device < 0 || device >= count) throw std::runtime_error("device ordinal " + std::to_string(device) + " out of range (" + std::to_string(count) + " visible)");
    c->device = device;
    G16_CUDA(cudaSetDevice(c->device));
    int prio_lo = 0, prio_hi = 0;
    G16_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    // the main stream carries computeH and then the H.Z MSM.  Measured (G16_TIMELINE=1): at high priority h is ready
    // after 6.8 ms instead of ~50, but the proof takes the same 69 ms -- the GPU is busy throughout and some MSM
    // has to finish last -- and the host-buffer entry point gets slower (75.9 vs 72.0 ms); default = normal priority
    const int prio_main = prio_lo;
    try {
        G16_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_main));
        for (auto& a : c->aux) G16_CUDA(cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking));
        for (auto& t : c->tails) G16_CUDA(cudaStreamCreateWithPriority(&t, cudaStreamNonBlocking, prio_hi));
        for (auto& e : c->ev) G16_CUDA(cudaEventCreate(&e));
    } catch (...) {
        ctx_destroy_single(c.release());
        throw;
    }
    return c.release();
}

void ctx_destroy_single(g16_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    for (auto& kv : ctx->pks) kv.second.release();
    for (auto& kv : ctx->r1cs) kv.second.release();
    for (auto& kv : ctx->bases) kv.second.table.release();
    for (auto& kv : ctx->commit_keys) { kv.second.basis.table.release(); kv.second.basis_exp_sigma.table.release(); }
    for (auto& kv : ctx->domains) kv.second.release();
    for (auto& w : ctx->ws) w.release();
    for (auto& s : ctx->stage) s.release();
    for (auto& s : ctx->file_stage) s.release();
    ctx->stager.release();
    ctx->fb_table_g1.release(); ctx->fb_table_g2.release();
    for (auto& e : ctx->ev) if (e) cudaEventDestroy(e);
    for (auto& a : ctx->aux) if (a) cudaStreamDestroy(a);
    for (auto& t : ctx->tails) if (t) cudaStreamDestroy(t);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

}  // namespace g16

extern "C" {

int g16_ctx_create(g16_ctx** out, int curve_id, const int* device_ids, int n_devices) {
    if (!out) return 1;
    *out = nullptr;
    if (curve_id != (G16_CURVE == 377 ? G16_CURVE_BLS12_377 : G16_CURVE_BN254)) return 3;   // one entry-point set per curve (curve_select.h)
    if (n_devices > 8 || (n_devices > 1 && (n_devices & (n_devices - 1)))) return 4;        // 1, 2, 4 or 8 GPUs
    if (n_devices >= 1 && !device_ids) return 1;
    g16_ctx* c = nullptr;
    try {
        c = ctx_create_single(n_devices >= 1 ? device_ids[0] : 0);
        if (n_devices > 1) c->multi = multi_create(c, device_ids, n_devices);
    } catch (const std::exception& e) {
        fprintf(stderr, "g16_ctx_create: %s\n", e.what());
        if (c) ctx_destroy_single(c);
        return 2;
    }
    *out = c;
    return 0;
}

int g16_device_count(void) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return 0; }
    return count;
}

void g16_ctx_destroy(g16_ctx* ctx) {
    if (!ctx) return;
    if (ctx->multi) { multi_destroy(ctx->multi); ctx->multi = nullptr; }
    ctx_destroy_single(ctx);
}

const char* g16_last_error(const g16_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int g16_ctx_set(g16_ctx* ctx, const char* key, int64_t value) {
    if (!ctx || !key) return 1;
    std::string k(key);
    if (k == "msm_c") ctx->msm_c = (int)value;
    else if (k == "msm_waves") ctx->msm_waves = (int)value;
    else if (k == "streams") ctx->streams = (int)value;
    else if (k == "msm_precompute") ctx->msm_precompute = (int)value;
    else if (k == "msm_pre_c") ctx->msm_pre_c = (int)value;
    else if (k == "msm_precompute_g1") ctx->msm_precompute_g1 = (int)value;
    else if (k == "dist_sync") ctx->dist_sync = (int)value;
    else if (k == "g2_weight") ctx->g2_weight = (int)value;
    else if (k == "stage_threads") ctx->stage_threads = (int)value;
    else if (k == "msm_ba_rounds") { for (auto& w : ctx->ws) w.ba_rounds = (int)value; }
    else if (k == "msm_ba_rounds_g1") { for (int i = 0; i < 4; i++) ctx->ws[i].ba_rounds = (int)value; }
    else if (k == "msm_ba_rounds_g2") ctx->ws[4].ba_rounds = (int)value;
    else { ctx->err = "unknown tunable " + k; return 1; }
    if (ctx->multi) multi_set(ctx->multi, key, value);
    return 0;
}

uint64_t g16_kernel_launch_count(void) { return launch_counter().load(); }
void* g16_ctx_stream(g16_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

}  // extern "C"

// =================================================================== proving key
namespace g16 {

void pk_validate(const g16_pk_desc* d) {
    if (!d) throw std::runtime_error("null argument");
    if (!is_pow2(d->domain_n)) throw std::runtime_error("domain_n must be a power of two");
    if (d->n_b2 != d->n_b) throw std::runtime_error("G2.B and G1.B must have the same length");
    if (d->n_z + 1 != d->domain_n && d->n_z != d->domain_n) throw std::runtime_error("G1.Z must have domain_n - 1 entries");
    if (d->n_k > d->n_wires) throw std::runtime_error("G1.K longer than the wire vector");
    if (d->n_wires >= 0xffffffffull) throw std::runtime_error("wire vector too long for 32-bit wire indices");
}

// slice shard_index of shard_count of every base array; Z follows the block distribution of h over the domain (so a
// distributed computeH leaves each rank's block of h exactly under its Z slice); the last block loses the dropped
// degree-(n-1) coefficient (h[:n-1], prove.go:235-237)
PkRanges pk_shard_ranges(const g16_pk_desc* d) {
    const uint32_t sc = d->shard_count ? d->shard_count : 1, si = d->shard_index;
    if (si >= sc) throw std::runtime_error("shard_index out of range");
    PkRanges rg;
    slice(d->n_a, si, sc, rg.lo[G16_OUT_AR], rg.hi[G16_OUT_AR]);
    slice(d->n_b, si, sc, rg.lo[G16_OUT_BS1], rg.hi[G16_OUT_BS1]);
    slice(d->n_k, si, sc, rg.lo[G16_OUT_KRS_K], rg.hi[G16_OUT_KRS_K]);
    slice(d->n_b, si, sc, rg.lo[G16_OUT_BS2], rg.hi[G16_OUT_BS2]);
    const uint64_t nz = d->n_z < d->domain_n ? d->n_z : d->domain_n - 1;
    uint64_t lo, hi;
    slice(d->domain_n, si, sc, lo, hi);
    rg.lo[G16_OUT_KRS_HZ] = lo > nz ? nz : lo;
    rg.hi[G16_OUT_KRS_HZ] = hi > nz ? nz : hi;
    return rg;
}

// src: G1.A, G1.B, G1.K, G1.Z, G2.B
void pk_upload_ranges(g16_ctx* ctx, const g16_pk_desc* d, const Src* src, const PkRanges& rg, uint64_t handle) {
    pk_validate(d);
    const uint64_t lens[5] = {d->n_a, d->n_b, d->n_k, d->n_z < d->domain_n ? d->n_z : d->domain_n - 1, d->n_b2};
    for (int i = 0; i < 5; i++)
        if (rg.lo[i] > rg.hi[i] || rg.hi[i] > lens[i]) throw std::runtime_error("proving-key slice out of range");
    cudaStream_t st = ctx->stream;
    PkDev pk;
    pk.domain_n = d->domain_n;
    pk.logn = ilog2_ceil(d->domain_n);
    pk.n_wires = d->n_wires;

    // wire index of every A / B base: position among the non-infinity wires (prove.go:152-178)
    std::vector<uint32_t> ma, mb, mk;
    ma.reserve(d->n_a); mb.reserve(d->n_b);
    for (uint64_t i = 0; i < d->n_wires; i++) {
        if (!d->infinity_a || !d->infinity_a[i]) ma.push_back((uint32_t)i);
        if (!d->infinity_b || !d->infinity_b[i]) mb.push_back((uint32_t)i);
    }
    if (ma.size() != d->n_a) throw std::runtime_error("InfinityA does not match len(G1.A): " + std::to_string(ma.size()) + " vs " + std::to_string(d->n_a));
    if (mb.size() != d->n_b) throw std::runtime_error("InfinityB does not match len(G1.B): " + std::to_string(mb.size()) + " vs " + std::to_string(d->n_b));
    mk.resize(d->n_k);
    for (uint64_t i = 0; i < d->n_k; i++) {
        mk[i] = d->k_index ? d->k_index[i] : (uint32_t)(d->n_wires - d->n_k + i);
        if (mk[i] >= d->n_wires) throw std::runtime_error("k_index out of range");
    }

    pk.ranges = rg;
    pk.w_lo = d->n_wires; pk.w_hi = 0;
    auto touch = [&](const std::vector<uint32_t>& m, uint64_t a, uint64_t b) {   // maps are not assumed monotone (k_index)
        for (uint64_t i = a; i < b; i++) { if (m[i] < pk.w_lo) pk.w_lo = m[i]; if ((uint64_t)m[i] + 1 > pk.w_hi) pk.w_hi = (uint64_t)m[i] + 1; }
    };
    try {
        uint64_t lo = rg.lo[G16_OUT_AR], hi = rg.hi[G16_OUT_AR];
        touch(ma, lo, hi);
        pk.n_a = hi - lo;
        upload_bases<G1>(ctx, pk.g1_a, src[0].at(lo * G1SZ), pk.n_a, pk.pre_c[G16_OUT_AR], st);
        upload(pk.map_a, ma.data() + lo, pk.n_a * 4, st);
        lo = rg.lo[G16_OUT_BS1]; hi = rg.hi[G16_OUT_BS1];
        touch(mb, lo, hi);
        pk.n_b = hi - lo;
        upload_bases<G1>(ctx, pk.g1_b, src[1].at(lo * G1SZ), pk.n_b, pk.pre_c[G16_OUT_BS1], st);
        upload(pk.map_b, mb.data() + lo, pk.n_b * 4, st);
        pk.same_b = rg.lo[G16_OUT_BS2] == lo && rg.hi[G16_OUT_BS2] == hi;
        lo = rg.lo[G16_OUT_BS2]; hi = rg.hi[G16_OUT_BS2];
        touch(mb, lo, hi);
        pk.n_b2 = hi - lo;
        upload_bases<G2>(ctx, pk.g2_b, src[4].at(lo * G2SZ), pk.n_b2, pk.pre_c[G16_OUT_BS2], st);
        if (!pk.same_b) upload(pk.map_b2, mb.data() + lo, pk.n_b2 * 4, st);
        lo = rg.lo[G16_OUT_KRS_K]; hi = rg.hi[G16_OUT_KRS_K];
        touch(mk, lo, hi);
        if (pk.w_hi <= pk.w_lo) pk.w_lo = pk.w_hi = 0;
        pk.n_k = hi - lo;
        // G1.K usually covers the wires of G1.A from some wire on (all private wires against all wires with a non-zero A
        // polynomial): padded with that many points at infinity in front, its array lines up with G1.A position by
        // position, and its MSM reuses G1.A's digits and bucket sort (the way bs1 reuses bs2's) instead of sorting the same
        // scalars a second time.  Whole-array uploads only (shards and the ranks of a group cut A and K differently).
        pk.k_pad = 0;
        {
            const uint64_t la = rg.lo[G16_OUT_AR], ha = rg.hi[G16_OUT_AR];
            const uint64_t pad = (ha - la) - (hi - lo);
            bool tail = la == 0 && ha == d->n_a && lo == 0 && hi == d->n_k && d->n_k >= PRECOMPUTE_MIN_POINTS && d->n_a >= d->n_k && pad <= 64;
            for (uint64_t i = 0; tail && i < d->n_k; i++) tail = mk[i] == ma[pad + i];
            if (tail) pk.k_pad = pad;
        }
        upload_bases<G1>(ctx, pk.g1_k, src[2].at(lo * G1SZ), pk.n_k, pk.pre_c[G16_OUT_KRS_K], st, pk.k_pad);
        if (pk.k_pad) { pk.n_k += pk.k_pad; pk.same_ak = true; upload(pk.map_k, ma.data(), pk.n_k * 4, st); }
        else { pk.same_ak = false; upload(pk.map_k, mk.data() + lo, pk.n_k * 4, st); }
        lo = rg.lo[G16_OUT_KRS_HZ]; hi = rg.hi[G16_OUT_KRS_HZ];
        pk.n_z = hi - lo;
        pk.z_off = lo;
        upload_bases<G1>(ctx, pk.g1_z, src[3].at(lo * G1SZ), pk.n_z, pk.pre_c[G16_OUT_KRS_HZ], st);

        // twiddles of the domain (replaces icicle.go:44-75)
        NttDomain& dom = ctx->domain(pk.logn);
        dom.table(TW_INV, st);
        dom.table(TW_COSET_FWD, st);
        dom.table(TW_COSET_INV, st);
        G16_CUDA(cudaStreamSynchronize(st));   // host vectors above go out of scope
    } catch (...) {
        cudaStreamSynchronize(st);
        pk.release();                          // nothing of a half-uploaded key stays allocated
        throw;
    }
    ctx->pks[handle] = pk;
}

}  // namespace g16

namespace {
int pk_upload_impl(g16_ctx* ctx, const g16_pk_desc* d, const Src* src, uint64_t* pk_handle) {
    return guarded(ctx, [&] {
        if (!d || !pk_handle) throw std::runtime_error("null argument");
        if (ctx->multi) { multi_pk_upload(ctx, d, src, pk_handle); return; }
        pk_validate(d);
        const uint64_t h = ctx->next_pk++;
        pk_upload_ranges(ctx, d, src, pk_shard_ranges(d), h);
        *pk_handle = h;
    });
}
}  // namespace

extern "C" int g16_pk_upload(g16_ctx* ctx, const g16_pk_desc* d, uint64_t* pk_handle) {
    if (!ctx) return 1;
    if (!d) { ctx->err = "null argument"; return 2; }
    const Src src[5] = {host_src(d->g1_a), host_src(d->g1_b), host_src(d->g1_k), host_src(d->g1_z), host_src(d->g2_b)};
    return pk_upload_impl(ctx, d, src, pk_handle);
}

// the point slices of a ProvingKey.WriteDump image (marshal.go:386-456) are raw memory, i.e.
// already the device format; they go file -> pinned staging -> HBM (sharded: only this shard's byte ranges are read).
extern "C" int g16_pk_upload_from_file(g16_ctx* ctx, const g16_pk_file_desc* f, uint64_t* pk_handle) {
    if (!ctx) return 1;
    if (!f || f->fd < 0) { ctx->err = "null argument / bad file descriptor"; return 2; }
    Src src[5];
    const uint64_t off[5] = {f->off_g1_a, f->off_g1_b, f->off_g1_k, f->off_g1_z, f->off_g2_b};
    for (int i = 0; i < 5; i++) { src[i].fd = f->fd; src[i].off = off[i]; }
    return pk_upload_impl(ctx, &f->key, src, pk_handle);
}

extern "C" int g16_pk_wire_range(g16_ctx* ctx, uint64_t pk_handle, uint64_t* lo, uint64_t* hi) {
    return guarded(ctx, [&] {
        if (!lo || !hi) throw std::runtime_error("null argument");
        if (ctx->multi) { multi_pk_wire_range(ctx, pk_handle, lo, hi); return; }
        auto it = ctx->pks.find(pk_handle);
        if (it == ctx->pks.end()) throw std::runtime_error("unknown proving-key handle");
        *lo = it->second.w_lo; *hi = it->second.w_hi;
    });
}

extern "C" int g16_pk_placement(g16_ctx* ctx, uint64_t pk_handle, int rank, uint64_t lo[5], uint64_t hi[5]) {
    return guarded(ctx, [&] {
        if (!lo || !hi) throw std::runtime_error("null argument");
        const PkDev* pk;
        if (ctx->multi) pk = &multi_pk_of_rank(ctx, pk_handle, rank);
        else {
            auto it = ctx->pks.find(pk_handle);
            if (it == ctx->pks.end()) throw std::runtime_error("unknown proving-key handle");
            if (rank != 0) throw std::runtime_error("rank out of range");
            pk = &it->second;
        }
        for (int i = 0; i < 5; i++) { lo[i] = pk->ranges.lo[i]; hi[i] = pk->ranges.hi[i]; }
    });
}

extern "C" int g16_pk_free(g16_ctx* ctx, uint64_t pk_handle) {
    return guarded(ctx, [&] {
        if (ctx->multi) { multi_pk_free(ctx, pk_handle); return; }
        auto it = ctx->pks.find(pk_handle);
        if (it == ctx->pks.end()) throw std::runtime_error("unknown proving-key handle");
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
        it->second.release();
        ctx->pks.erase(it);
    });
}

// =================================================================== prove
namespace g16 {

// phase 0: the whole device section.  phase 1: enqueue the four MSMs over W and return (they run while the caller
// produces h, e.g. the distributed computeH with its exchanges); phase 2: enqueue the H.Z MSM, wait for all five, finish.
void prove_device(g16_ctx* ctx, PkDev& pk, const Fr::T* W_d, Fr::T* a_d, Fr::T* b_d, Fr::T* c_d, void* out_points,
                  g16_stats* stats, uint64_t launches0, cudaEvent_t w_ready, bool h_given, int phase, ProvePartial* partial, cudaEvent_t acc_gate) {
    cudaStream_t st = ctx->stream;
    NttDomain& dom = ctx->domain(pk.logn);
    cudaEvent_t* ev = ctx->ev;
    MsmPlan plan[5];
    MsmStats ms[5];
    for (int i = 0; i < 5; i++) { ms[i].acc_begin = ev[16 + 2 * i]; ms[i].acc_end = ev[17 + 2 * i]; }
    // The four MSMs over W do not depend on h: with overlap on they run on their own streams next to computeH
    // (the reference serialises everything, icicle.go:396-410); the sort/reduce tails of one MSM then hide
    // under the accumulation of another.  Order of issue: G2 first (longest).
    const bool overlap = ctx->streams > 1;
    // precomputed tables are used unless the caller forces a window size ("msm_c") or switched them off after upload
    auto pre = [&](int i) { return (ctx->msm_c == 0 && ctx->msm_precompute) ? pk.pre_c[i] : 0; };
    cudaStream_t sA = overlap ? ctx->aux[0] : st, sB = overlap ? ctx->aux[1] : st, sK = overlap ? ctx->aux[2] : st,
                 s2 = overlap ? ctx->aux[3] : st;
    if (phase != 2) {
        G16_CUDA(cudaEventRecord(ev[0], st));
        if (overlap) {
            cudaEvent_t ready = w_ready ? w_ready : ev[0];
            for (int i = 0; i < 4; i++) G16_CUDA(cudaStreamWaitEvent(ctx->aux[i], ready, 0));
        }
    } else {
        for (int i = 0; i < 5; i++) plan[i] = ctx->pend_plan[i];
    }
    auto run = [&](int idx, cudaStream_t s, auto&& fn) {
        G16_CUDA(cudaEventRecord(ev[1 + idx], s));
        fn(s);
        G16_CUDA(cudaEventRecord(ev[6 + idx], s));
    };
    if (phase != 2) run(G16_OUT_BS2, s2, [&](cudaStream_t s) {
        msm_g2_enqueue(ctx->ws[G16_OUT_BS2], W_d, (pk.same_b ? pk.map_b : pk.map_b2).as<uint32_t>(), pk.g2_b.as<G2::Aff>(), pk.n_b2, ctx->msm_c, ctx->msm_waves, s, plan[4], &ms[4], overlap ? ctx->tails[4] : nullptr, pre(G16_OUT_BS2), nullptr, acc_gate); });
    if (phase != 1) {
        G16_CUDA(cudaEventRecord(ev[12], st));
        if (!h_given) compute_h_device(dom, a_d, b_d, c_d, st);   // h (bit-reversed) in a_d; h_given: a_d is this shard's block of h
        G16_CUDA(cudaEventRecord(ev[13], st));
    }
    if (phase != 2) run(G16_OUT_AR, sA, [&](cudaStream_t s) {
        msm_g1_enqueue(ctx->ws[G16_OUT_AR], W_d, pk.map_a.as<uint32_t>(), pk.g1_a.as<G1::Aff>(), pk.n_a, ctx->msm_c, ctx->msm_waves, s, plan[0], &ms[0], overlap ? ctx->tails[0] : nullptr, pre(G16_OUT_AR), nullptr, acc_gate); });
    if (phase != 1) run(G16_OUT_KRS_HZ, st, [&](cudaStream_t s) {
        msm_g1_enqueue(ctx->ws[G16_OUT_KRS_HZ], a_d + (h_given ? 0 : pk.z_off), nullptr, pk.g1_z.as<G1::Aff>(), pk.n_z, ctx->msm_c, ctx->msm_waves, s, plan[3], &ms[3], overlap ? ctx->tails[3] : nullptr, pre(G16_OUT_KRS_HZ)); });
    if (phase != 2) run(G16_OUT_BS1, sB, [&](cudaStream_t s) {
        msm_g1_enqueue(ctx->ws[G16_OUT_BS1], W_d, pk.map_b.as<uint32_t>(), pk.g1_b.as<G1::Aff>(), pk.n_b, ctx->msm_c, ctx->msm_waves, s, plan[1], &ms[1], overlap ? ctx->tails[1] : nullptr, pre(G16_OUT_BS1),
                       pk.same_b && pk.n_b == pk.n_b2 ? &ctx->ws[G16_OUT_BS2] : nullptr, acc_gate); });   // same scalars, same map: reuse the G2 MSM's sort
    if (phase != 2) run(G16_OUT_KRS_K, sK, [&](cudaStream_t s) {
        msm_g1_enqueue(ctx->ws[G16_OUT_KRS_K], W_d, pk.map_k.as<uint32_t>(), pk.g1_k.as<G1::Aff>(), pk.n_k, ctx->msm_c, ctx->msm_waves, s, plan[2], &ms[2], overlap ? ctx->tails[2] : nullptr, pre(G16_OUT_KRS_K),
                       pk.same_ak ? &ctx->ws[G16_OUT_AR] : nullptr, acc_gate); });   // same wires, same positions: reuse G1.A's sort
    if (phase == 1) {
        for (int i = 0; i < 5; i++) ctx->pend_plan[i] = plan[i];
        return;
    }
    if (overlap) {
        const int aux_of[4] = {G16_OUT_AR, G16_OUT_BS1, G16_OUT_KRS_K, G16_OUT_BS2};
        for (int i = 0; i < 4; i++) G16_CUDA(cudaStreamWaitEvent(st, ev[6 + aux_of[i]], 0));
    }
    G16_CUDA(cudaEventRecord(ev[14], st));
    G16_CUDA(cudaStreamSynchronize(st));
    auto t0 = std::chrono::steady_clock::now();
    char* out = (char*)out_points;
    const uint64_t ns[5] = {pk.n_a, pk.n_b, pk.n_k, pk.n_z, pk.n_b2};
    if (partial) {
        for (int i = 0; i < 4; i++) partial->g1[i] = ns[i] ? msm_g1_finish_xyzz(ctx->ws[i], plan[i]) : G1::inf();
        partial->g2 = ns[4] ? msm_g2_finish_xyzz(ctx->ws[4], plan[4]) : G2::inf();
    } else {
        for (int i = 0; i < 4; i++) copy_out(out + G1SZ * i, ns[i] ? msm_g1_finish(ctx->ws[i], plan[i]) : G1::aff_inf());
        copy_out(out + 4 * G1SZ, ns[4] ? msm_g2_finish(ctx->ws[4], plan[4]) : G2::aff_inf());
    }
    auto t1 = std::chrono::steady_clock::now();
    if (stats) {
        G16_CUDA(cudaEventElapsedTime(&stats->ms_compute_h, ev[12], ev[13]));
        for (int i = 0; i < 5; i++) {
            G16_CUDA(cudaEventElapsedTime(&stats->ms_msm[i], ev[1 + i], ev[6 + i]));
            stats->msm_c[i] = plan[i].c; stats->msm_windows[i] = plan[i].Wd; stats->msm_n[i] = ns[i] - (i == G16_OUT_KRS_K ? pk.k_pad : 0);
            if (ns[i]) G16_CUDA(cudaEventElapsedTime(&stats->ms_accumulate[i], ev[16 + 2 * i], ev[17 + 2 * i]));
        }
        G16_CUDA(cudaEventElapsedTime(&stats->ms_total, ev[0], ev[14]));
        stats->ms_host = std::chrono::duration<float, std::milli>(t1 - t0).count();
        stats->d2h_bytes = 0;
        for (int i = 0; i < 5; i++) stats->d2h_bytes += (uint64_t)(1 + plan[i].nfold) * plan[i].W * (i == 4 ? sizeof(G2::Xyzz) : sizeof(G1::Xyzz));
        stats->kernel_launches = (uint32_t)(launch_counter().load() - launches0);
    }
    static const bool timeline = getenv("G16_TIMELINE") != nullptr;   // diagnostic: event offsets from the start of the prove
    if (timeline) {
        auto at = [&](cudaEvent_t e) { float ms = 0; cudaEventElapsedTime(&ms, ev[0], e); return ms; };
        const char* names[5] = {"ar", "bs1", "krs_k", "krs_hz", "bs2"};
        fprintf(stderr, "[g16 timeline] computeH %.2f-%.2f  end %.2f\n", at(ev[12]), at(ev[13]), at(ev[14]));
        for (int i = 0; i < 5; i++)
            if (ns[i]) fprintf(stderr, "[g16 timeline] %-7s msm %.2f-%.2f  accumulate %.2f-%.2f\n", names[i], at(ev[1 + i]), at(ev[6 + i]),
                               at(ev[16 + 2 * i]), at(ev[17 + 2 * i]));
    }
}

}  // namespace g16

extern "C" int g16_prove_bn254_dev(g16_ctx* ctx, uint64_t pk_handle, const void* W_d, void* a_d, void* b_d, void* c_d,
                                   void* out_points, g16_stats* stats) {
    return guarded(ctx, [&] {
        if (!W_d || !a_d || !b_d || !c_d || !out_points) throw std::runtime_error("null argument");
        if (stats) memset(stats, 0, sizeof(*stats));
        if (ctx->multi) {   // device pointers live on the group's first GPU; every rank pulls its share over NVLink
            MultiIn in; in.W = W_d; in.a = a_d; in.b = b_d; in.c = c_d; in.device = true;
            multi_prove(ctx, pk_handle, in, out_points, stats);
            return;
        }
        auto it = ctx->pks.find(pk_handle);
        if (it == ctx->pks.end()) throw std::runtime_error("unknown proving-key handle");
        uint64_t l0 = launch_counter().load();
        prove_device(ctx, it->second, (const Fr::T*)W_d, (Fr::T*)a_d, (Fr::T*)b_d, (Fr::T*)c_d, out_points, stats, l0, nullptr);
    });
}

extern "C" int g16_prove_bn254(g16_ctx* ctx, uint64_t pk_handle, const void* W, uint64_t n_wires, const void* a,
                               const void* b, const void* c, uint64_t n_constraints, void* out_points, g16_stats* stats) {
    return guarded(ctx, [&] {
        if (!W || !a || !b || !c || !out_points) throw std::runtime_error("null argument");
        if (stats) memset(stats, 0, sizeof(*stats));
        if (ctx->multi) {
            MultiIn in; in.W = W; in.a = a; in.b = b; in.c = c; in.n_wires = n_wires; in.n_constraints = n_constraints;
            multi_prove(ctx, pk_handle, in, out_points, stats);
            return;
        }
        auto it = ctx->pks.find(pk_handle);
        if (it == ctx->pks.end()) throw std::runtime_error("unknown proving-key handle");
        PkDev& pk = it->second;
        if (n_wires != pk.n_wires) throw std::runtime_error("witness length does not match the proving key");
        if (n_constraints > pk.domain_n) throw std::runtime_error("more constraints than the domain holds");
        uint64_t l0 = launch_counter().load();
        cudaStream_t st = ctx->stream;
        const size_t nb = (size_t)pk.domain_n * 32, cb = (size_t)n_constraints * 32;
        G16_CUDA(cudaEventRecord(ctx->ev[26], st));
        // W first: the four MSMs over W start on their own streams while a, b, c are still in flight
        // a sharded key reads only the wires [w_lo, w_hi) of W: nothing else crosses PCIe
        ctx->stage[3].reserve((size_t)n_wires * 32);
        for (int i = 0; i < 3; i++) ctx->stage[i].reserve(nb);
        const size_t w_bytes = (size_t)(pk.w_hi - pk.w_lo) * 32;
        h2d_caller(ctx, (char*)ctx->stage[3].p + (size_t)pk.w_lo * 32, (const char*)W + (size_t)pk.w_lo * 32, w_bytes, st);
        G16_CUDA(cudaEventRecord(ctx->ev[15], st));
        // pageable a, b, c are staged by this thread (h2d_caller blocks while it gathers): enqueue the W MSMs first
        const bool split = cb >= ((size_t)1 << 20) && HostStager::pageable(a);
        if (split) prove_device(ctx, pk, ctx->stage[3].as<Fr::T>(), nullptr, nullptr, nullptr, nullptr, nullptr, l0, ctx->ev[15], false, 1);
        const void* src[3] = {a, b, c};
        for (int i = 0; i < 3; i++) {      // pad to the domain size (icicle.go:463-467)
            h2d_caller(ctx, ctx->stage[i].p, src[i], cb, st);
            if (nb > cb) G16_CUDA(cudaMemsetAsync((char*)ctx->stage[i].p + cb, 0, nb - cb, st));
        }
        G16_CUDA(cudaEventRecord(ctx->ev[27], st));
        prove_device(ctx, pk, ctx->stage[3].as<Fr::T>(), ctx->stage[0].as<Fr::T>(), ctx->stage[1].as<Fr::T>(),
                     ctx->stage[2].as<Fr::T>(), out_points, stats, l0, ctx->ev[15], false, split ? 2 : 0);
        if (stats) {
            G16_CUDA(cudaEventElapsedTime(&stats->ms_h2d, ctx->ev[26], ctx->ev[27]));
            G16_CUDA(cudaEventElapsedTime(&stats->ms_total, ctx->ev[26], ctx->ev[14]));
            stats->h2d_bytes = 3 * cb + (uint64_t)w_bytes;
        }
    });
}

// =================================================================== resident bases
namespace g16 {
void bases_upload_one(g16_ctx* ctx, const void* bases, uint64_t n, bool g2, bool from_device, int src_device, uint64_t handle) {
    g16_ctx::Bases b;
    b.n = n; b.g2 = g2;
    cudaStream_t st = ctx->stream;
    try {
        if (from_device && n) {   // upload_bases copies host -> device; a device source is copied into row 0 here
            int pre_c = 0;
            const size_t row = (size_t)n * (g2 ? G2SZ : G1SZ);
            if (ctx->msm_precompute && n >= PRECOMPUTE_MIN_POINTS) {
                const int c = msm_pre_choose_c(n, g2, ctx->msm_pre_c);
                size_t free_b = 0, total_b = 0;
                G16_CUDA(cudaMemGetInfo(&free_b, &total_b));
                if (row * (size_t)(255 / c + 1) <= free_b / 2) pre_c = c;
            }
            b.table.reserve(pre_c ? row * (size_t)(255 / pre_c + 1) : row);
            if (src_device == ctx->device) G16_CUDA(cudaMemcpyAsync(b.table.p, bases, row, cudaMemcpyDeviceToDevice, st));
            else G16_CUDA(cudaMemcpyPeerAsync(b.table.p, ctx->device, bases, src_device, row, st));
            if (pre_c) { if (g2) msm_g2_precompute(b.table.as<G2::Aff>(), n, pre_c, st); else msm_g1_precompute(b.table.as<G1::Aff>(), n, pre_c, st); }
            b.pre_c = pre_c;
        } else if (g2) {
            upload_bases<G2>(ctx, b.table, bases, n, b.pre_c, st);
        } else {
            upload_bases<G1>(ctx, b.table, bases, n, b.pre_c, st);
        }
        G16_CUDA(cudaStreamSynchronize(st));
    } catch (...) {
        b.table.release();
        throw;
    }
    ctx->bases[handle] = b;
}
}  // namespace g16

// a group context shards arrays of at least this many points per GPU over its ranks; smaller ones stay on the first GPU
static constexpr uint64_t GROUP_MIN_POINTS_PER_RANK = 1 << 16;

extern "C" int g16_bases_upload(g16_ctx* ctx, const void* bases, uint64_t n, int g2, int from_device, uint64_t* handle) {
    return guarded(ctx, [&] {
        if (!handle || (n && !bases)) throw std::runtime_error("null argument");
        if (ctx->multi && n >= GROUP_MIN_POINTS_PER_RANK * 8) { multi_bases_upload(ctx, bases, n, g2, from_device, handle); return; }
        const uint64_t h = ctx->next_pk++;
        bases_upload_one(ctx, bases, n, g2 != 0, from_device != 0, ctx->device, h);
        *handle = h;
    });
}
extern "C" int g16_bases_free(g16_ctx* ctx, uint64_t handle) {
    return guarded(ctx, [&] {
        if (multi_has_bases(ctx, handle)) { multi_bases_free(ctx, handle); return; }
        auto it = ctx->bases.find(handle);
        if (it == ctx->bases.end()) throw std::runtime_error("unknown bases handle");
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
        it->second.table.release();
        ctx->bases.erase(it);
    });
}
extern "C" int g16_msm_resident_dev(g16_ctx* ctx, uint64_t handle, const void* scalars_d, void* out) {
    return guarded(ctx, [&] {
        if (multi_has_bases(ctx, handle)) { if (!out) throw std::runtime_error("null argument"); multi_msm_resident(ctx, handle, scalars_d, true, out); return; }
        auto it = ctx->bases.find(handle);
        if (it == ctx->bases.end()) throw std::runtime_error("unknown bases handle");
        if (!out || (it->second.n && !scalars_d)) throw std::runtime_error("null argument");
        const g16_ctx::Bases& b = it->second;
        const int pre_c = (ctx->msm_c == 0 && ctx->msm_precompute) ? b.pre_c : 0;
        MsmPlan pl;
        if (b.g2) {
            if (b.n) { msm_g2_enqueue(ctx->ws[4], (const Fr::T*)scalars_d, nullptr, b.table.as<G2::Aff>(), b.n, ctx->msm_c, ctx->msm_waves, ctx->stream, pl, nullptr, nullptr, pre_c);
                       G16_CUDA(cudaStreamSynchronize(ctx->stream)); }
            copy_out(out, b.n ? msm_g2_finish(ctx->ws[4], pl) : G2::aff_inf());
        } else {
            if (b.n) { msm_g1_enqueue(ctx->ws[0], (const Fr::T*)scalars_d, nullptr, b.table.as<G1::Aff>(), b.n, ctx->msm_c, ctx->msm_waves, ctx->stream, pl, nullptr, nullptr, pre_c);
                       G16_CUDA(cudaStreamSynchronize(ctx->stream)); }
            copy_out(out, b.n ? msm_g1_finish(ctx->ws[0], pl) : G1::aff_inf());
        }
    });
}

extern "C" int g16_msm_resident(g16_ctx* ctx, uint64_t handle, const void* scalars, void* out) {
    // host-scalar form of g16_msm_resident_dev: one staged H2D copy, then the same path
    if (!ctx) return 1;
    if (multi_has_bases(ctx, handle)) return guarded(ctx, [&] { if (!out) throw std::runtime_error("null argument"); multi_msm_resident(ctx, handle, scalars, false, out); });
    uint64_t n = 0;
    int rc = guarded(ctx, [&] {
        auto it = ctx->bases.find(handle);
        if (it == ctx->bases.end()) throw std::runtime_error("unknown bases handle");
        n = it->second.n;
        if (n && !scalars) throw std::runtime_error("null argument");
        upload(ctx->stage[3], scalars, (size_t)n * 32, ctx->stream);
    });
    return rc ? rc : g16_msm_resident_dev(ctx, handle, ctx->stage[3].p, out);
}

// =================================================================== BSB22 commitment keys
// pedersen.ProvingKey{Basis, BasisExpSigma} (setup.go:57,295-305) resident on the device.  g16_commit is what the
// bsb22 hint (prove.go:94: CommitmentKeys[i].Commit) and the proof of knowledge (prove.go:124: ProveKnowledge)
// become: both are MSMs of the same committed values, so one call uploads the values once and returns both points.
extern "C" int g16_commitment_key_upload(g16_ctx* ctx, const void* basis, const void* basis_exp_sigma, uint64_t n,
                                         uint64_t* key_handle) {
    return guarded(ctx, [&] {
        if (!key_handle || (n && (!basis || !basis_exp_sigma))) throw std::runtime_error("null argument");
        g16_ctx::CommitKey k;
        cudaStream_t st = ctx->stream;
        k.basis.n = k.basis_exp_sigma.n = n;
        try {
            upload_bases<G1>(ctx, k.basis.table, basis, n, k.basis.pre_c, st);
            upload_bases<G1>(ctx, k.basis_exp_sigma.table, basis_exp_sigma, n, k.basis_exp_sigma.pre_c, st);
            G16_CUDA(cudaStreamSynchronize(st));
        } catch (...) {
            k.basis.table.release(); k.basis_exp_sigma.table.release();
            throw;
        }
        uint64_t h = ctx->next_pk++;
        ctx->commit_keys[h] = k;
        *key_handle = h;
    });
}
extern "C" int g16_commitment_key_free(g16_ctx* ctx, uint64_t key_handle) {
    return guarded(ctx, [&] {
        auto it = ctx->commit_keys.find(key_handle);
        if (it == ctx->commit_keys.end()) throw std::runtime_error("unknown commitment-key handle");
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
        it->second.basis.table.release();
        it->second.basis_exp_sigma.table.release();
        ctx->commit_keys.erase(it);
    });
}
extern "C" int g16_commit(g16_ctx* ctx, uint64_t key_handle, const void* values, uint64_t n_values,
                          void* commitment_out, void* pok_out) {
    return guarded(ctx, [&] {
        auto it = ctx->commit_keys.find(key_handle);
        if (it == ctx->commit_keys.end()) throw std::runtime_error("unknown commitment-key handle");
        const g16_ctx::CommitKey& k = it->second;
        // pedersen.Commit / ProveKnowledge: "must have as many values as basis elements"
        if (n_values != k.basis.n) throw std::runtime_error("must have as many values as basis elements");
        if (n_values && !values) throw std::runtime_error("null argument");
        if (!commitment_out && !pok_out) throw std::runtime_error("no output requested");
        void* outs[2] = {commitment_out, pok_out};
        if (!n_values) {   // MultiExp over nothing: the point at infinity
            for (void* o : outs) if (o) copy_out(o, G1::aff_inf());
            return;
        }
        upload(ctx->stage[3], values, (size_t)n_values * 32, ctx->stream);
        G16_CUDA(cudaEventRecord(ctx->ev[15], ctx->stream));
        const g16_ctx::Bases* bs[2] = {&k.basis, &k.basis_exp_sigma};
        cudaStream_t sts[2] = {ctx->stream, ctx->aux[0]};
        MsmPlan pl[2];
        for (int j = 0; j < 2; j++) {   // the two MSMs side by side: own workspace and stream each
            if (!outs[j]) continue;
            if (j) G16_CUDA(cudaStreamWaitEvent(sts[j], ctx->ev[15], 0));
            const int pre_c = (ctx->msm_c == 0 && ctx->msm_precompute) ? bs[j]->pre_c : 0;
            msm_g1_enqueue(ctx->ws[j], ctx->stage[3].as<Fr::T>(), nullptr, bs[j]->table.as<G1::Aff>(), n_values, ctx->msm_c,
                           ctx->msm_waves, sts[j], pl[j], nullptr, nullptr, pre_c);
        }
        for (int j = 0; j < 2; j++) {
            if (!outs[j]) continue;
            G16_CUDA(cudaStreamSynchronize(sts[j]));
            copy_out(outs[j], msm_g1_finish(ctx->ws[j], pl[j]));
        }
    });
}

// =================================================================== a, b, c on the device
extern "C" int g16_r1cs_upload(g16_ctx* ctx, const g16_r1cs_desc* d, uint64_t* handle) {
    return guarded(ctx, [&] {
        if (!d || !handle) throw std::runtime_error("null argument");
        if (d->n_constraints >= 0xffffffffull || d->n_wires >= 0xffffffffull) throw std::runtime_error("constraint system too large for 32-bit indices");
        if (d->n_coeffs && !d->coeffs) throw std::runtime_error("null coefficient table");
        R1csDev r;
        r.n_constraints = d->n_constraints; r.n_wires = d->n_wires; r.n_coeffs = d->n_coeffs;
        cudaStream_t st = ctx->stream;
        try {
            upload(r.coeffs, d->coeffs, (size_t)d->n_coeffs * 32, st);
            for (int m = 0; m < 3; m++) {
                const uint32_t* rp = d->row_ptr[m];
                if (!rp) throw std::runtime_error("null row_ptr");
                const uint64_t nnz = rp[d->n_constraints];
                if (rp[0] != 0) throw std::runtime_error("row_ptr must start at 0");
                for (uint64_t i = 0; i < d->n_constraints; i++)
                    if (rp[i] > rp[i + 1]) throw std::runtime_error("row_ptr is not monotone");
                if (nnz && (!d->wire[m] || !d->coeff_id[m])) throw std::runtime_error("null term arrays");
                for (uint64_t t = 0; t < nnz; t++) {
                    if (d->wire[m][t] >= d->n_wires) throw std::runtime_error("term wire id out of range");
                    if (d->coeff_id[m][t] >= d->n_coeffs) throw std::runtime_error("term coefficient id out of range");
                }
                upload(r.row_ptr[m], rp, (size_t)(d->n_constraints + 1) * 4, st);
                upload(r.wire[m], d->wire[m], (size_t)nnz * 4, st);
                upload(r.cid[m], d->coeff_id[m], (size_t)nnz * 4, st);
            }
            G16_CUDA(cudaStreamSynchronize(st));
        } catch (...) {
            r.release();
            throw;
        }
        uint64_t h = ctx->next_pk++;
        ctx->r1cs[h] = r;
        if (ctx->multi) {
            try { multi_r1cs_upload(ctx, d, h); }
            catch (...) { ctx->r1cs[h].release(); ctx->r1cs.erase(h); throw; }
        }
        *handle = h;
    });
}

extern "C" int g16_r1cs_free(g16_ctx* ctx, uint64_t handle) {
    return guarded(ctx, [&] {
        auto it = ctx->r1cs.find(handle);
        if (it == ctx->r1cs.end()) throw std::runtime_error("unknown R1CS handle");
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
        it->second.release();
        ctx->r1cs.erase(it);
        if (ctx->multi) multi_r1cs_free(ctx, handle);
    });
}

extern "C" int g16_r1cs_eval_dev(g16_ctx* ctx, uint64_t handle, const void* W_d, void* a_d, void* b_d, void* c_d, uint64_t domain_n) {
    return guarded(ctx, [&] {
        auto it = ctx->r1cs.find(handle);
        if (it == ctx->r1cs.end()) throw std::runtime_error("unknown R1CS handle");
        if (!W_d || !a_d || !b_d || !c_d) throw std::runtime_error("null argument");
        if (domain_n < it->second.n_constraints) throw std::runtime_error("domain smaller than the constraint count");
        r1cs_eval(it->second, (const Fr::T*)W_d, (Fr::T*)a_d, (Fr::T*)b_d, (Fr::T*)c_d, domain_n, ctx->stream);
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

extern "C" int g16_prove_bn254_wires(g16_ctx* ctx, uint64_t pk_handle, uint64_t r1cs_handle, const void* W, uint64_t n_wires,
                                     void* out_points, g16_stats* stats) {
    return guarded(ctx, [&] {
        if (!W || !out_points) throw std::runtime_error("null argument");
        if (stats) memset(stats, 0, sizeof(*stats));
        if (ctx->multi) {
            MultiIn in; in.W = W; in.n_wires = n_wires; in.r1cs = r1cs_handle;
            multi_prove(ctx, pk_handle, in, out_points, stats);
            return;
        }
        auto it = ctx->pks.find(pk_handle);
        if (it == ctx->pks.end()) throw std::runtime_error("unknown proving-key handle");
        auto ir = ctx->r1cs.find(r1cs_handle);
        if (ir == ctx->r1cs.end()) throw std::runtime_error("unknown R1CS handle");
        PkDev& pk = it->second;
        const R1csDev& r = ir->second;
        if (n_wires != pk.n_wires || n_wires != r.n_wires) throw std::runtime_error("witness length does not match the key / constraint system");
        if (r.n_constraints > pk.domain_n) throw std::runtime_error("more constraints than the domain holds");
        uint64_t l0 = launch_counter().load();
        cudaStream_t st = ctx->stream;
        const size_t nb = (size_t)pk.domain_n * 32;
        G16_CUDA(cudaEventRecord(ctx->ev[26], st));
        // the WHOLE wire vector is uploaded, also for a sharded key: the constraint matrices gather arbitrary wires
        // (a public wire may appear in O only, i.e. in none of this shard's A / B / K maps)
        ctx->stage[3].reserve((size_t)n_wires * 32);
        h2d_caller(ctx, ctx->stage[3].p, W, (size_t)n_wires * 32, st);
        G16_CUDA(cudaEventRecord(ctx->ev[15], st));
        G16_CUDA(cudaEventRecord(ctx->ev[27], st));
        for (int i = 0; i < 3; i++) ctx->stage[i].reserve(nb);
        r1cs_eval(r, ctx->stage[3].as<Fr::T>(), ctx->stage[0].as<Fr::T>(), ctx->stage[1].as<Fr::T>(), ctx->stage[2].as<Fr::T>(), pk.domain_n, st);
        prove_device(ctx, pk, ctx->stage[3].as<Fr::T>(), ctx->stage[0].as<Fr::T>(), ctx->stage[1].as<Fr::T>(),
                     ctx->stage[2].as<Fr::T>(), out_points, stats, l0, ctx->ev[15]);
        if (stats) {
            G16_CUDA(cudaEventElapsedTime(&stats->ms_h2d, ctx->ev[26], ctx->ev[27]));
            G16_CUDA(cudaEventElapsedTime(&stats->ms_total, ctx->ev[26], ctx->ev[14]));
            stats->h2d_bytes = (uint64_t)n_wires * 32;
        }
    });
}

// =================================================================== primitives
namespace {

void msm_g1_dev_impl(g16_ctx* ctx, const void* s_d, const void* b_d, uint64_t n, void* out) {
    MsmPlan pl;
    if (n) {
        msm_g1_enqueue(ctx->ws[0], (const Fr::T*)s_d, nullptr, (const G1::Aff*)b_d, n, ctx->msm_c, ctx->msm_waves, ctx->stream, pl, nullptr);
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    copy_out(out, n ? msm_g1_finish(ctx->ws[0], pl) : G1::aff_inf());
}
void msm_g2_dev_impl(g16_ctx* ctx, const void* s_d, const void* b_d, uint64_t n, void* out) {
    MsmPlan pl;
    if (n) {
        msm_g2_enqueue(ctx->ws[4], (const Fr::T*)s_d, nullptr, (const G2::Aff*)b_d, n, ctx->msm_c, ctx->msm_waves, ctx->stream, pl, nullptr);
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    copy_out(out, n ? msm_g2_finish(ctx->ws[4], pl) : G2::aff_inf());
}

// gnark fft.Domain semantics on a device array (see include/g16b200.h)
void ntt_dev_impl(g16_ctx* ctx, Fr::T* x, uint64_t n, int inverse, int coset, int decimation) {
    if (!is_pow2(n)) throw std::runtime_error("NTT size must be a power of two");
    int logn = ilog2_ceil(n);
    NttDomain& d = ctx->domain(logn);
    cudaStream_t st = ctx->stream;
    NttBatch b{{x, x, x}, 1};
    if (n == 1) return;   // every variant is the identity on one element (g^0 = 1, 1/n = 1)
    if (!inverse) {
        if (decimation == G16_DIT) {
            ntt_dit(d, b, coset ? TW_COSET_FWD : TW_FWD, st);
        } else {
            if (coset) fr_scale_powers(x, n, logn, Fr::one(), d.coset_gen, 0, 1, st);   // a[i] *= g^i (natural index)
            ntt_dif(d, b, TW_FWD, st);
        }
    } else {
        if (decimation == G16_DIF) {
            ntt_dif(d, b, coset ? TW_COSET_INV : TW_INV, st);
            fr_scale_powers(x, n, logn, d.n_inv, Fr::one(), 0, 0, st);
        } else {
            ntt_dit(d, b, TW_INV, st);
            if (coset) fr_scale_powers(x, n, logn, d.n_inv, d.coset_inv, 0, 1, st);   // natural order out
            else fr_scale_powers(x, n, logn, d.n_inv, Fr::one(), 0, 0, st);
        }
    }
}

}  // namespace

extern "C" {

int g16_msm_g1_dev(g16_ctx* ctx, const void* s_d, const void* b_d, uint64_t n, void* out) {
    return guarded(ctx, [&] { msm_g1_dev_impl(ctx, s_d, b_d, n, out); });
}
int g16_msm_g2_dev(g16_ctx* ctx, const void* s_d, const void* b_d, uint64_t n, void* out) {
    return guarded(ctx, [&] { msm_g2_dev_impl(ctx, s_d, b_d, n, out); });
}
int g16_msm_g1(g16_ctx* ctx, const void* scalars, const void* bases, uint64_t n, void* out) {
    return guarded(ctx, [&] {
        upload(ctx->stage[0], scalars, n * 32, ctx->stream);
        upload(ctx->stage[1], bases, n * G1SZ, ctx->stream);
        msm_g1_dev_impl(ctx, ctx->stage[0].p, ctx->stage[1].p, n, out);
    });
}
int g16_msm_g2(g16_ctx* ctx, const void* scalars, const void* bases, uint64_t n, void* out) {
    return guarded(ctx, [&] {
        upload(ctx->stage[0], scalars, n * 32, ctx->stream);
        upload(ctx->stage[1], bases, n * G2SZ, ctx->stream);
        msm_g2_dev_impl(ctx, ctx->stage[0].p, ctx->stage[1].p, n, out);
    });
}

int g16_ntt_dev(g16_ctx* ctx, void* data_d, uint64_t n, int inverse, int coset, int decimation) {
    return guarded(ctx, [&] {
        // a group context distributes the transform shapes computeH is made of (one exchange); the rest runs on its first GPU
        if (multi_can_distribute(ctx, n) && n >= GROUP_MIN_POINTS_PER_RANK * 8 && multi_ntt_dev(ctx, data_d, n, inverse, coset, decimation)) return;
        ntt_dev_impl(ctx, (Fr::T*)data_d, n, inverse, coset, decimation);
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}
int g16_ntt(g16_ctx* ctx, void* data, uint64_t n, int inverse, int coset, int decimation) {
    return guarded(ctx, [&] {
        upload(ctx->stage[0], data, n * 32, ctx->stream);
        ntt_dev_impl(ctx, ctx->stage[0].as<Fr::T>(), n, inverse, coset, decimation);
        G16_CUDA(cudaMemcpyAsync(data, ctx->stage[0].p, n * 32, cudaMemcpyDeviceToHost, ctx->stream));
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

int g16_compute_h_dev(g16_ctx* ctx, void* a_d, void* b_d, void* c_d, uint64_t domain_n) {
    return guarded(ctx, [&] {
        if (!is_pow2(domain_n)) throw std::runtime_error("domain_n must be a power of two");
        if (multi_can_distribute(ctx, domain_n) && domain_n >= GROUP_MIN_POINTS_PER_RANK * 8) { multi_compute_h_dev(ctx, a_d, b_d, c_d, domain_n); return; }
        compute_h_device(ctx->domain(ilog2_ceil(domain_n)), (Fr::T*)a_d, (Fr::T*)b_d, (Fr::T*)c_d, ctx->stream);
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}
int g16_compute_h(g16_ctx* ctx, const void* a, const void* b, const void* c, uint64_t n_constraints, uint64_t domain_n, void* h_out) {
    return guarded(ctx, [&] {
        if (!is_pow2(domain_n) || n_constraints > domain_n) throw std::runtime_error("bad domain size");
        const size_t nb = (size_t)domain_n * 32, cb = (size_t)n_constraints * 32;
        const void* src[3] = {a, b, c};
        for (int i = 0; i < 3; i++) {
            ctx->stage[i].reserve(nb);
            if (cb) G16_CUDA(cudaMemcpyAsync(ctx->stage[i].p, src[i], cb, cudaMemcpyHostToDevice, ctx->stream));
            if (nb > cb) G16_CUDA(cudaMemsetAsync((char*)ctx->stage[i].p + cb, 0, nb - cb, ctx->stream));
        }
        compute_h_device(ctx->domain(ilog2_ceil(domain_n)), ctx->stage[0].as<Fr::T>(), ctx->stage[1].as<Fr::T>(), ctx->stage[2].as<Fr::T>(), ctx->stream);
        G16_CUDA(cudaMemcpyAsync(h_out, ctx->stage[0].p, nb, cudaMemcpyDeviceToHost, ctx->stream));
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

// No counterpart in the reference: computeH for callers that may not allocate device memory (see ntt.cu).  The only
// scratch is the library's static device array; this call performs no cudaMalloc / cudaFree.
int g16_compute_h_noalloc_dev(g16_ctx* ctx, const void* a_d, const void* b_d, const void* c_d, uint64_t n_constraints,
                              uint64_t domain_n, void* h_out_d) {
    return guarded(ctx, [&] {
        if (!h_out_d || (n_constraints && (!a_d || !b_d || !c_d))) throw std::runtime_error("null argument");
        compute_h_noalloc(ctx->device, (const Fr::T*)a_d, (const Fr::T*)b_d, (const Fr::T*)c_d, n_constraints, domain_n,
                          (Fr::T*)h_out_d, ctx->stream);
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

int g16_fr_mul_vec(g16_ctx* ctx, const void* a, const void* b, uint64_t n, void* out) {
    return guarded(ctx, [&] {
        upload(ctx->stage[0], a, n * 32, ctx->stream);
        upload(ctx->stage[1], b, n * 32, ctx->stream);
        fr_mul_vec(ctx->stage[0].as<Fr::T>(), ctx->stage[0].as<Fr::T>(), ctx->stage[1].as<Fr::T>(), n, ctx->stream);
        if (n) G16_CUDA(cudaMemcpyAsync(out, ctx->stage[0].p, n * 32, cudaMemcpyDeviceToHost, ctx->stream));
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

int g16_g1_fixed_base_mul_dev(g16_ctx* ctx, const void* s_d, uint64_t n, void* out_d) {
    return guarded(ctx, [&] {
        fixed_base_g1(ctx->fb_table_g1, (const Fr::T*)s_d, n, (G1::Aff*)out_d, ctx->stream);
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}
int g16_g2_fixed_base_mul_dev(g16_ctx* ctx, const void* s_d, uint64_t n, void* out_d) {
    return guarded(ctx, [&] {
        fixed_base_g2(ctx->fb_table_g2, (const Fr::T*)s_d, n, (G2::Aff*)out_d, ctx->stream);
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}
int g16_g1_fixed_base_mul(g16_ctx* ctx, const void* scalars, uint64_t n, void* out) {
    return guarded(ctx, [&] {
        upload(ctx->stage[0], scalars, n * 32, ctx->stream);
        ctx->stage[1].reserve(n * G1SZ + 16);
        fixed_base_g1(ctx->fb_table_g1, ctx->stage[0].as<Fr::T>(), n, ctx->stage[1].as<G1::Aff>(), ctx->stream);
        if (n) G16_CUDA(cudaMemcpyAsync(out, ctx->stage[1].p, n * G1SZ, cudaMemcpyDeviceToHost, ctx->stream));
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}
int g16_g2_fixed_base_mul(g16_ctx* ctx, const void* scalars, uint64_t n, void* out) {
    return guarded(ctx, [&] {
        upload(ctx->stage[0], scalars, n * 32, ctx->stream);
        ctx->stage[1].reserve(n * G2SZ + 16);
        fixed_base_g2(ctx->fb_table_g2, ctx->stage[0].as<Fr::T>(), n, ctx->stage[1].as<G2::Aff>(), ctx->stream);
        if (n) G16_CUDA(cudaMemcpyAsync(out, ctx->stage[1].p, n * G2SZ, cudaMemcpyDeviceToHost, ctx->stream));
        G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

}  // extern "C"

// =================================================================== host-side group helpers + epilogue
namespace {

template <class C>
typename C::Aff aff_add(const typename C::Aff& a, const typename C::Aff& b) {
    typename C::Xyzz t = C::from_aff(a);
    C::madd(t, b);
    return C::to_aff(t);
}
template <class C>
typename C::Aff aff_mul(const typename C::Aff& a, const Fr::T& k_mont) {
    Fr::T k = Fr::from_mont(k_mont);
    return C::to_aff(C::mul(C::from_aff(a), k.l, 8));
}
template <class T>
T rd(const void* p) { T v; memcpy(&v, p, sizeof(T)); return v; }

}  // namespace

extern "C" {

int g16_g1_add(const void* a, const void* b, void* out) {
    if (!a || !b || !out) return 1;
    copy_out(out, aff_add<G1>(rd<G1::Aff>(a), rd<G1::Aff>(b)));
    return 0;
}
int g16_g2_add(const void* a, const void* b, void* out) {
    if (!a || !b || !out) return 1;
    copy_out(out, aff_add<G2>(rd<G2::Aff>(a), rd<G2::Aff>(b)));
    return 0;
}
int g16_g1_scalar_mul(const void* a, const void* k, void* out) {
    if (!a || !k || !out) return 1;
    copy_out(out, aff_mul<G1>(rd<G1::Aff>(a), rd<Fr::T>(k)));
    return 0;
}
int g16_g2_scalar_mul(const void* a, const void* k, void* out) {
    if (!a || !k || !out) return 1;
    copy_out(out, aff_mul<G2>(rd<G2::Aff>(a), rd<Fr::T>(k)));
    return 0;
}

// prove.go:180-195 (kr = -r*s, deltas), :209-224 (ar), :202-212 (bs1), :251-279 (krs), :297-302 (Bs)
int g16_assemble_proof_bn254(const void* pk_points, const void* r_, const void* s_, const void* msm_points, void* proof_out) {
    if (!pk_points || !r_ || !s_ || !msm_points || !proof_out) return 1;
    const char* pp = (const char*)pk_points;
    const char* mp = (const char*)msm_points;
    G1::Aff alpha = rd<G1::Aff>(pp), beta = rd<G1::Aff>(pp + G1SZ), delta = rd<G1::Aff>(pp + 2 * G1SZ);
    G2::Aff beta2 = rd<G2::Aff>(pp + 3 * G1SZ), delta2 = rd<G2::Aff>(pp + 3 * G1SZ + G2SZ);
    Fr::T r = rd<Fr::T>(r_), s = rd<Fr::T>(s_);
    Fr::T kr = Fr::neg(Fr::mul(r, s));
    G1::Aff d_r = aff_mul<G1>(delta, r), d_s = aff_mul<G1>(delta, s), d_kr = aff_mul<G1>(delta, kr);
    G1::Aff ar = aff_add<G1>(aff_add<G1>(rd<G1::Aff>(mp + G1SZ * G16_OUT_AR), alpha), d_r);
    G1::Aff bs1 = aff_add<G1>(aff_add<G1>(rd<G1::Aff>(mp + G1SZ * G16_OUT_BS1), beta), d_s);
    G1::Aff krs = aff_add<G1>(rd<G1::Aff>(mp + G1SZ * G16_OUT_KRS_K), d_kr);
    krs = aff_add<G1>(krs, rd<G1::Aff>(mp + G1SZ * G16_OUT_KRS_HZ));
    krs = aff_add<G1>(krs, aff_mul<G1>(ar, s));
    krs = aff_add<G1>(krs, aff_mul<G1>(bs1, r));
    G2::Aff bs = aff_add<G2>(aff_add<G2>(rd<G2::Aff>(mp + 4 * G1SZ), aff_mul<G2>(delta2, s)), beta2);
    char* o = (char*)proof_out;
    copy_out(o, ar);
    copy_out(o + G1SZ, bs);
    copy_out(o + G1SZ + G2SZ, krs);
    return 0;
}

}  // extern "C"

// =================================================================== integer-pipe peak microbenchmark
namespace {

template <bool WIDE>
__global__ void __launch_bounds__(256) k_imad_peak(uint32_t* sink, uint32_t a0, uint32_t b0, int iters) {
    uint32_t a = a0 + threadIdx.x, b = b0 | 1u;
    if (WIDE) {
        uint64_t x0 = threadIdx.x, x1 = a, x2 = b, x3 = 7, x4 = 11, x5 = 13, x6 = 17, x7 = 19;
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 8; u++) {
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x0) : "r"(a), "r"(b));
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x1) : "r"(a), "r"(b));
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x2) : "r"(a), "r"(b));
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x3) : "r"(a), "r"(b));
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x4) : "r"(a), "r"(b));
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x5) : "r"(a), "r"(b));
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x6) : "r"(a), "r"(b));
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x7) : "r"(a), "r"(b));
            }
        }
        uint64_t x = x0 ^ x1 ^ x2 ^ x3 ^ x4 ^ x5 ^ x6 ^ x7;
        if (x == 0x123456789abcdefull) sink[0] = (uint32_t)x;
    } else {
        uint32_t x0 = threadIdx.x, x1 = a, x2 = b, x3 = 7, x4 = 11, x5 = 13, x6 = 17, x7 = 19;
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int u = 0; u < 8; u++) {
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x0) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x1) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x2) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x3) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x4) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x5) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x6) : "r"(a), "r"(b));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x7) : "r"(a), "r"(b));
            }
        }
        uint32_t x = x0 ^ x1 ^ x2 ^ x3 ^ x4 ^ x5 ^ x6 ^ x7;
        if (x == 0x12345678u) sink[0] = x;
    }
}

template <bool WIDE>
double imad_rate(g16_ctx* ctx, uint32_t* sink) {
    int sms = 0;
    G16_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
    const int iters = 4096, blocks = sms * 8;
    cudaStream_t st = ctx->stream;
    k_imad_peak<WIDE><<<blocks, 256, 0, st>>>(sink, 3, 5, 64);   // warm-up
    G16_LAUNCH_CHECK();
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        G16_CUDA(cudaEventRecord(ctx->ev[10], st));
        k_imad_peak<WIDE><<<blocks, 256, 0, st>>>(sink, 3, 5, iters);
        G16_LAUNCH_CHECK();
        G16_CUDA(cudaEventRecord(ctx->ev[11], st));
        G16_CUDA(cudaStreamSynchronize(st));
        float ms = 0;
        G16_CUDA(cudaEventElapsedTime(&ms, ctx->ev[10], ctx->ev[11]));
        if (ms < best) best = ms;
    }
    double ops = (double)blocks * 256 * iters * 64;
    return ops / (best * 1e-3);
}

}  // namespace

extern "C" int g16_imad_peak(g16_ctx* ctx, double* imad_per_s, double* imad_wide_per_s) {
    return guarded(ctx, [&] {
        ctx->stage[0].reserve(64);
        if (imad_per_s) *imad_per_s = imad_rate<false>(ctx, ctx->stage[0].as<uint32_t>());
        if (imad_wide_per_s) *imad_wide_per_s = imad_rate<true>(ctx, ctx->stage[0].as<uint32_t>());
    });
}

// =================================================================== pipe-rate microbenchmarks (roofline denominators)
namespace {

// independent carry chains shaped like Field::cmad_n: 4 x (mad.lo.cc, madc.hi.cc) + addc per chain, 4 chains in flight
__global__ void __launch_bounds__(256) k_wide_carry_peak(uint32_t* sink, uint32_t a0, int iters) {
#ifdef __CUDA_ARCH__
    uint32_t a[4] = {a0 + threadIdx.x, a0 * 3 + 1, a0 * 5 + 2, a0 * 7 + 3};
    uint32_t acc[4][9];
    for (int c = 0; c < 4; c++)
        for (int k = 0; k < 9; k++) acc[c][k] = threadIdx.x + c * 9 + k;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < 4; c++) {
            uint32_t bi = acc[c][0] | 1u;
            acc[c][0] = ptx::mad_lo_cc(a[0], bi, acc[c][0]);
            acc[c][1] = ptx::madc_hi_cc(a[0], bi, acc[c][1]);
#pragma unroll
            for (int j = 2; j < 8; j += 2) {
                acc[c][j] = ptx::madc_lo_cc(a[j >> 1], bi, acc[c][j]);
                acc[c][j + 1] = ptx::madc_hi_cc(a[j >> 1], bi, acc[c][j + 1]);
            }
            acc[c][8] = ptx::addc(acc[c][8], 0);
        }
    }
    uint32_t x = 0;
    for (int c = 0; c < 4; c++)
        for (int k = 0; k < 9; k++) x ^= acc[c][k];
    if (x == 0x12345678u) sink[0] = x;
#endif
}

// 4 independent Montgomery multiplications per iteration, all in registers
template <class F>
__global__ void __launch_bounds__(256) k_fr_mul_peak(uint32_t* sink, uint32_t seed, int iters) {
    typename F::T x[4], y = F::one();
    y.l[0] += seed + threadIdx.x;
    for (int c = 0; c < 4; c++) { x[c] = F::r2(); x[c].l[0] += c + threadIdx.x; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < 4; c++) x[c] = F::mul(x[c], y);
    }
    uint32_t v = 0;
    for (int c = 0; c < 4; c++)
        for (int k = 0; k < F::N; k++) v ^= x[c].l[k];
    if (v == 0x12345678u) sink[0] = v;
}

template <class K>
double rate_of(g16_ctx* ctx, K launch, double ops_per_launch) {
    cudaStream_t st = ctx->stream;
    launch(8);
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        G16_CUDA(cudaEventRecord(ctx->ev[10], st));
        launch(0);
        G16_CUDA(cudaEventRecord(ctx->ev[11], st));
        G16_CUDA(cudaStreamSynchronize(st));
        float ms = 0;
        G16_CUDA(cudaEventElapsedTime(&ms, ctx->ev[10], ctx->ev[11]));
        if (ms < best) best = ms;
    }
    return ops_per_launch / (best * 1e-3);
}

}  // namespace

// out[0] = IMAD.WIDE.U32.X (carry chain) issue rate per second; out[1] / out[2] = in-register Fr / Fp multiplications per second
extern "C" int g16_pipe_rates(g16_ctx* ctx, double* out2) {
    return guarded(ctx, [&] {
        if (!out2) throw std::runtime_error("null argument");
        ctx->stage[0].reserve(64);
        uint32_t* sink = ctx->stage[0].as<uint32_t>();
        int sms = 0;
        G16_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
        const int blocks = sms * 8, it1 = 2048, it2 = 512;
        cudaStream_t st = ctx->stream;
        out2[0] = rate_of(ctx, [&](int small) { k_wide_carry_peak<<<blocks, 256, 0, st>>>(sink, 3, small ? small : it1); G16_LAUNCH_CHECK(); },
                          (double)blocks * 256 * it1 * 16);
        out2[1] = rate_of(ctx, [&](int small) { k_fr_mul_peak<Fr><<<blocks, 256, 0, st>>>(sink, 3, small ? small : it2); G16_LAUNCH_CHECK(); },
                          (double)blocks * 256 * it2 * 4);
        out2[2] = rate_of(ctx, [&](int small) { k_fr_mul_peak<Fp><<<blocks, 256, 0, st>>>(sink, 3, small ? small : it2); G16_LAUNCH_CHECK(); },
                          (double)blocks * 256 * it2 * 4);
    });
}


// =================================================================== distributed computeH building blocks + MSM-only prove
extern "C" {

int g16_dist_ntt_stages_dev(g16_ctx* ctx, void* data_d, int count, uint64_t n_global, int world, int rank, int table, int dif,
                            int cyclic, int split) {
    return guarded(ctx, [&] {
        if (!is_pow2(n_global) || world < 1 || (world & (world - 1)) || count < 1 || count > 3 || table < 0 || table > 3)
            throw std::runtime_error("g16_dist_ntt_stages_dev: bad arguments");
        int g = ilog2_ceil((uint64_t)world);
        NttDomain& d = ctx->domain(ilog2_ceil(n_global));
        uint64_t n_local = n_global >> g;
        Fr::T* p = (Fr::T*)data_d;
        NttBatch b{{p, p + (count > 1 ? n_local : 0), p + (count > 2 ? 2 * n_local : 0)}, count};
        ntt_dist_stages(d, b, (NttTable)table, dif != 0, g, (uint32_t)rank, cyclic != 0, split, ctx->stream);
        if (ctx->dist_sync) G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}
int g16_dist_permute_dev(g16_ctx* ctx, const void* src_d, void* dst_d, int count, uint64_t n_local, int world, int block_layout, int to_wire) {
    return guarded(ctx, [&] {
        if (world < 1 || n_local % (uint64_t)world) throw std::runtime_error("g16_dist_permute_dev: n_local must be a multiple of world");
        ntt_dist_permute((const Fr::T*)src_d, (Fr::T*)dst_d, count, n_local, (uint32_t)world, block_layout, to_wire, ctx->stream);
        if (ctx->dist_sync) G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}
int g16_dist_pointwise_h_dev(g16_ctx* ctx, void* abc_d, uint64_t n_local, uint64_t n_global) {
    return guarded(ctx, [&] {
        if (!is_pow2(n_global)) throw std::runtime_error("bad domain size");
        NttDomain& d = ctx->domain(ilog2_ceil(n_global));
        Fr::T* p = (Fr::T*)abc_d;
        fr_pointwise_h(p, p + n_local, p + 2 * n_local, n_local, d.h_c1, d.h_c2, ctx->stream);
        if (ctx->dist_sync) G16_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}
int g16_prove_msms_dev(g16_ctx* ctx, uint64_t pk_handle, const void* W_d, const void* h_local_d, void* out_points, g16_stats* stats) {
    return guarded(ctx, [&] {
        auto it = ctx->pks.find(pk_handle);
        if (it == ctx->pks.end()) throw std::runtime_error("unknown proving-key handle");
        if (!W_d || !h_local_d || !out_points) throw std::runtime_error("null argument");
        uint64_t l0 = launch_counter().load();
        if (stats) memset(stats, 0, sizeof(*stats));
        const bool pending = ctx->pend_pk == pk_handle;
        if (ctx->pend_pk && !pending) throw std::runtime_error("g16_prove_msms_begin_dev is pending for another proving key");
        ctx->pend_pk = 0;
        prove_device(ctx, it->second, (const Fr::T*)W_d, (Fr::T*)h_local_d, nullptr, nullptr, out_points, stats, l0, nullptr, true, pending ? 2 : 0);
    });
}
int g16_prove_msms_begin_dev(g16_ctx* ctx, uint64_t pk_handle, const void* W_d) {
    return guarded(ctx, [&] {
        auto it = ctx->pks.find(pk_handle);
        if (it == ctx->pks.end()) throw std::runtime_error("unknown proving-key handle");
        if (!W_d) throw std::runtime_error("null argument");
        if (ctx->pend_pk) throw std::runtime_error("g16_prove_msms_begin_dev: a previous begin has not been finished by g16_prove_msms_dev");
        prove_device(ctx, it->second, (const Fr::T*)W_d, nullptr, nullptr, nullptr, nullptr, nullptr, 0, nullptr, true, 1);
        ctx->pend_pk = pk_handle;
    });
}

}  // extern "C"
