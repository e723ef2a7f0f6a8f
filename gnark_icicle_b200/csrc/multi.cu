// Several GPUs behind ONE context: g16_ctx_create(..., device_ids, n) with n in {2, 4, 8} builds a group of single-GPU
// contexts ("ranks", one persistent host thread each) and g16_pk_upload / g16_prove_bn254 then shard internally, so a
// host that can only make the reference's one in-process call (icicle_bn254.Prove, backend/groth16/bn254/icicle/
// icicle.go:133) reaches every GPU of the box.  No NCCL and no second process: the exchanges of the
// distributed computeH are stores into peer memory over NVLink / NVSwitch issued by the kernel that changes the layout
// (ntt.cu k_dist_exchange), ordered by CUDA events across the devices.
//
//  * key placement (SURVEY.md 8(e) levels 1 + 2): the four MSMs over W are laid end to end on a cost line (a G2 point
//    costs g2_weight/100 G1 points) and the line is cut into n equal parts, so every GPU holds one or two LARGE pieces
//    (G2.B alone fills ~3 of 8 GPUs) instead of 1/n of all four: larger windows, fewer sort heads and tails; G1.Z is cut
//    by the block distribution of h, so every GPU multiplies the block of h that computeH leaves on it, last;
//  * computeH (level 3): every size-N transform is "local stages - exchange - local stages" between the distribution
//    D(rb) (rank = position bits [rb, rb+g)) and the block distribution; a, b, c are loaded from the host straight into
//    D(rb) (strided copies of 2^rb x 32 B), h leaves block-distributed in bit-reversed order and the GPUs that hold a
//    slice of G1.Z pull the matching range of h from its owners;
//  * the W MSMs start as soon as W is on the device and run beside computeH; the five partial results of every rank
//    are summed on the host in XYZZ form and go affine once.
#include "ctx.cuh"

#include <chrono>

extern "C" int g16_prove_bn254(g16_ctx*, uint64_t, const void*, uint64_t, const void*, const void*, const void*, uint64_t, void*, g16_stats*);
extern "C" int g16_prove_bn254_dev(g16_ctx*, uint64_t, const void*, void*, void*, void*, void*, g16_stats*);
extern "C" int g16_prove_bn254_wires(g16_ctx*, uint64_t, uint64_t, const void*, uint64_t, void*, g16_stats*);

namespace g16 {

namespace {

struct RankState {
    g16_ctx* ctx = nullptr;
    cudaStream_t hs = nullptr;       // the distributed computeH (high priority: other GPUs wait for its exchanges)
    DevBuf buf[2];                   // ping-pong arrays of the distributed computeH, 3 x n/world Fr each
    DevBuf hbuf;                     // the range of h under this rank's G1.Z slice when other ranks own (part of) it
    cudaEvent_t ev_x[4] = {};        // [k]: this rank's exchange k has landed everywhere; [3]: its block of h is final
    cudaEvent_t ev_in = nullptr;     // rank 0: the caller's device-resident inputs are ready
    cudaEvent_t ev_w = nullptr;      // this rank's 1/n slice of the wire vector is on its GPU
    cudaEvent_t mark[12] = {};       // G16_TIMELINE: phases of the distributed computeH on hs
    ProvePartial part;
    g16_stats st;
};

struct GroupPk {
    int active = 1;                  // ranks that hold a slice (1: a domain too small to distribute lives on rank 0)
    uint64_t domain_n = 0, n_wires = 0;
};

}  // namespace

struct GroupBases { uint64_t n = 0; bool g2 = false; };

struct MultiGpu {
    int n = 0, g = 0;
    std::vector<RankState> rk;
    Team team;
    HostBarrier bar;
    std::map<uint64_t, GroupPk> pks;
    std::map<uint64_t, GroupBases> bases;
};

namespace {

int split_stage(int L, int g) {
    int x = L / 2;
    if (x > L - g) x = L - g;
    if (x < g) x = g;
    return x;
}

void set_device(const RankState& R) { G16_CUDA(cudaSetDevice(R.ctx->device)); }

// dst[arr][q] = src[arr][i(q)], i(q) = the q-th position of this rank under D(rb): device-resident inputs on another
// (or the same) GPU are read through the peer mapping
struct PullArgs { const Fr::T* src[3]; Fr::T* dst; uint64_t n_local; int count, rb, g; uint32_t rank; };
__global__ void __launch_bounds__(256) k_dist_pull(PullArgs a) {
    const uint64_t total = (uint64_t)a.count * a.n_local, rmask = (1ull << a.rb) - 1;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t arr = e / a.n_local, q = e % a.n_local;
        const uint64_t i = ((q >> a.rb) << (a.rb + a.g)) | ((uint64_t)a.rank << a.rb) | (q & rmask);
        const uint4* s = reinterpret_cast<const uint4*>(a.src[arr] + i);
        uint4* d = reinterpret_cast<uint4*>(a.dst + e);
        const uint4 lo = s[0], hi = s[1];
        d[0] = lo; d[1] = hi;
    }
}

// one array of a, b, c from host memory into D(rb): rows of 2^rb elements, `n_constraints` elements exist, the rest is zero
void load_host_rows(g16_ctx* c, int world, Fr::T* dst, const void* src, uint64_t n_constraints, uint64_t n, uint64_t n_local, int rb, int g, uint32_t rank, cudaStream_t st) {
    const uint64_t row = 1ull << rb, rows = n_local >> rb, pitch = row << g;
    auto start = [&](uint64_t hi) { return hi * pitch + (uint64_t)rank * row; };
    if (n_constraints < n) G16_CUDA(cudaMemsetAsync(dst, 0, (size_t)n_local * sizeof(Fr::T), st));
    uint64_t full = 0;                                   // rows that lie wholly below n_constraints
    if (n_constraints >= (uint64_t)rank * row + row) {
        full = (n_constraints - (uint64_t)rank * row - row) / pitch + 1;
        if (full > rows) full = rows;
    }
    const char* s = (const char*)src;
    const bool staged = (size_t)full * row * 32 >= ((size_t)1 << 20) && HostStager::pageable(src);
    if (staged && !c->stager.threads) {
        int t = c->stage_threads;
        if (t <= 0) { t = (int)std::thread::hardware_concurrency() / (2 * world); t = t < 2 ? 2 : (t > 8 ? 8 : t); }
        c->stager.start(t);
    }
    if (full) {
        if (staged) c->stager.copy_rows(dst, s + start(0) * 32, row * 32, pitch * 32, full, st);
        else if (g == 0 || rows == 1) G16_CUDA(cudaMemcpyAsync(dst, s + start(0) * 32, (size_t)full * row * 32, cudaMemcpyHostToDevice, st));
        else G16_CUDA(cudaMemcpy2DAsync(dst, row * 32, s + start(0) * 32, pitch * 32, row * 32, full, cudaMemcpyHostToDevice, st));
    }
    if (full < rows && start(full) < n_constraints)
        G16_CUDA(cudaMemcpyAsync(dst + full * row, s + start(full) * 32, (size_t)(n_constraints - start(full)) * 32, cudaMemcpyHostToDevice, st));
}

// Placement of the key over `parts` GPUs.  The four MSMs over W are laid end to end on a cost line [BS2 | AR | KRS_K | BS1]
// (a G2 point costs g2_weight G1 points) and the line is cut into equal parts: every GPU gets one or two LARGE pieces.
// G1.Z is cut by the block distribution of h instead: every GPU multiplies its own block of h, which computeH leaves in
// place, and does so last -- the W pieces keep it busy until h exists (with Z on the cost line the GPUs that held only Z
// sat idle for the whole of computeH: 3.7 of 13.5 ms at N = 8).
std::vector<PkRanges> place_key(const uint64_t lens[5], int parts, double g2_weight, uint64_t domain_n) {
    const int order[4] = {G16_OUT_BS2, G16_OUT_AR, G16_OUT_KRS_K, G16_OUT_BS1};
    double w[5] = {1, 1, 1, 1, 1}, begin[5] = {0, 0, 0, 0, 0}, total = 0;
    for (int k = 0; k < 4; k++) {
        const int i = order[k];
        w[i] = i == G16_OUT_BS2 ? g2_weight : 1.0;
        begin[i] = total;
        total += (double)lens[i] * w[i];
    }
    // position of cut c (cost c * total / parts) inside array i, snapped to the array's ends when it falls within 3 % of a part
    auto cut = [&](int c, int i) -> uint64_t {
        if (c <= 0) return 0;
        if (c >= parts) return lens[i];
        const double cost = total * c / parts, tol = 0.03 * total / parts, len = (double)lens[i] * w[i];
        if (cost <= begin[i] + tol) return 0;
        if (cost >= begin[i] + len - tol) return lens[i];
        uint64_t p = (uint64_t)((cost - begin[i]) / w[i] + 0.5);
        return p > lens[i] ? lens[i] : p;
    };
    std::vector<PkRanges> out(parts);
    const uint64_t nz = lens[G16_OUT_KRS_HZ];
    for (int r = 0; r < parts; r++) {
        for (int k = 0; k < 4; k++) { const int i = order[k]; out[r].lo[i] = cut(r, i); out[r].hi[i] = cut(r + 1, i); }
        uint64_t lo, hi;
        slice(domain_n, (uint32_t)r, (uint32_t)parts, lo, hi);
        out[r].lo[G16_OUT_KRS_HZ] = lo > nz ? nz : lo;
        out[r].hi[G16_OUT_KRS_HZ] = hi > nz ? nz : hi;
    }
    return out;
}

void upload_r1cs_to(g16_ctx* c, const g16_r1cs_desc* d, uint64_t handle) {
    R1csDev r;
    r.n_constraints = d->n_constraints; r.n_wires = d->n_wires; r.n_coeffs = d->n_coeffs;
    cudaStream_t st = c->stream;
    try {
        upload(r.coeffs, d->coeffs, (size_t)d->n_coeffs * 32, st);
        for (int m = 0; m < 3; m++) {
            const uint64_t nnz = d->row_ptr[m][d->n_constraints];
            upload(r.row_ptr[m], d->row_ptr[m], (size_t)(d->n_constraints + 1) * 4, st);
            upload(r.wire[m], d->wire[m], (size_t)nnz * 4, st);
            upload(r.cid[m], d->coeff_id[m], (size_t)nnz * 4, st);
        }
        G16_CUDA(cudaStreamSynchronize(st));
    } catch (...) {
        r.release();
        throw;
    }
    c->r1cs[handle] = r;
}

}  // namespace

// =================================================================== group life cycle
MultiGpu* multi_create(g16_ctx* rank0, const int* device_ids, int n_devices) {
    std::unique_ptr<MultiGpu> mg(new MultiGpu());
    mg->n = n_devices;
    mg->g = ilog2_ceil((uint64_t)n_devices);
    mg->rk.resize(n_devices);
    try {
        mg->rk[0].ctx = rank0;
        for (int r = 1; r < n_devices; r++) mg->rk[r].ctx = ctx_create_single(device_ids[r]);
        // every GPU of the group maps every other one (NVLink / NVSwitch peer access); the same device may be listed
        // several times (ranks emulated on one GPU: tests), which needs no mapping
        for (int a = 0; a < n_devices; a++) {
            RankState& R = mg->rk[a];
            set_device(R);
            for (int b = 0; b < n_devices; b++) {
                const int da = R.ctx->device, db = mg->rk[b].ctx->device;
                if (da == db) continue;
                int can = 0;
                G16_CUDA(cudaDeviceCanAccessPeer(&can, da, db));
                if (!can) throw std::runtime_error("GPU " + std::to_string(da) + " cannot map GPU " + std::to_string(db) + " (no peer access)");
                cudaError_t e = cudaDeviceEnablePeerAccess(db, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else G16_CUDA(e);
            }
            int prio_lo = 0, prio_hi = 0;
            G16_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
            G16_CUDA(cudaStreamCreateWithPriority(&R.hs, cudaStreamNonBlocking, prio_hi));
            for (auto& e : R.ev_x) G16_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            G16_CUDA(cudaEventCreateWithFlags(&R.ev_in, cudaEventDisableTiming));
            G16_CUDA(cudaEventCreateWithFlags(&R.ev_w, cudaEventDisableTiming));
            for (auto& e : R.mark) G16_CUDA(cudaEventCreate(&e));
        }
        G16_CUDA(cudaSetDevice(rank0->device));
    } catch (...) {
        multi_destroy(mg.release());
        cudaSetDevice(rank0->device);
        throw;
    }
    mg->team.start(n_devices);
    return mg.release();
}

void multi_destroy(MultiGpu* mg) {
    if (!mg) return;
    mg->team.shutdown();
    for (int r = 0; r < mg->n; r++) {
        RankState& R = mg->rk[r];
        if (!R.ctx) continue;
        cudaSetDevice(R.ctx->device);
        cudaDeviceSynchronize();
        R.buf[0].release(); R.buf[1].release(); R.hbuf.release();
        for (auto& e : R.ev_x) if (e) cudaEventDestroy(e);
        if (R.ev_in) cudaEventDestroy(R.ev_in);
        if (R.ev_w) cudaEventDestroy(R.ev_w);
        for (auto& e : R.mark) if (e) cudaEventDestroy(e);
        if (R.hs) cudaStreamDestroy(R.hs);
        if (r > 0) ctx_destroy_single(R.ctx);   // rank 0 belongs to the caller (g16_ctx_destroy)
    }
    cudaSetDevice(mg->rk[0].ctx->device);
    delete mg;
}

void multi_set(MultiGpu* mg, const char* key, int64_t value) {
    const std::string k(key);
    for (int r = 1; r < mg->n; r++) {
        g16_ctx* c = mg->rk[r].ctx;
        if (k == "msm_c") c->msm_c = (int)value;
        else if (k == "msm_waves") c->msm_waves = (int)value;
        else if (k == "streams") c->streams = (int)value;
        else if (k == "msm_precompute") c->msm_precompute = (int)value;
        else if (k == "msm_pre_c") c->msm_pre_c = (int)value;
        else if (k == "msm_precompute_g1") c->msm_precompute_g1 = (int)value;
        else if (k == "msm_ba_rounds") { for (auto& w : c->ws) w.ba_rounds = (int)value; }
        else if (k == "msm_ba_rounds_g1") { for (int i = 0; i < 4; i++) c->ws[i].ba_rounds = (int)value; }
        else if (k == "msm_ba_rounds_g2") c->ws[4].ba_rounds = (int)value;
    }
}

// =================================================================== key
void multi_pk_upload(g16_ctx* ctx, const g16_pk_desc* d, const Src* src, uint64_t* pk_handle) {
    MultiGpu& mg = *ctx->multi;
    pk_validate(d);
    if (d->shard_count > 1) throw std::runtime_error("a multi-GPU context shards the key itself: shard_count must be 0 or 1");
    GroupPk gp;
    gp.domain_n = d->domain_n; gp.n_wires = d->n_wires;
    const int L = ilog2_ceil(d->domain_n);
    gp.active = L >= 2 * mg.g + 4 ? mg.n : 1;   // every rank needs a few stages on either side of the exchanges
    const uint64_t lens[5] = {d->n_a, d->n_b, d->n_k, d->n_z < d->domain_n ? d->n_z : d->domain_n - 1, d->n_b2};
    // cost of a G2 point in G1 points: "g2_weight" if set, else by group size -- the smaller the pieces, the more a G2 piece
    // pays for its smaller window and its exposed bucket reduction (measured per piece at N = 2 / 4 / 8: ~3.0 / 3.4 / 4.0
    // before the G2 rounds moved their inversions to Fp, which took 9 % off a G2 MSM: swept again at N = 2, best at 2.75;
    // the other two scaled by the same factor, then by what the 17-bit tables of small G2 pieces gained; per-piece times
    // from tools/prof_piece.py: N = 4: 1.6 M G2 points ~12.7 ms against 5 M G1 points ~13.2; N = 8: 0.73 M ~7.6 against 2.7 M ~8.1)
    const double w2 = ctx->g2_weight > 0 ? ctx->g2_weight / 100.0 : (gp.active >= 8 ? 3.5 : gp.active >= 4 ? 3.0 : 2.75);
    std::vector<PkRanges> place = place_key(lens, gp.active, w2, d->domain_n);
    PkRanges none;
    for (int i = 0; i < 5; i++) none.lo[i] = none.hi[i] = 0;
    const uint64_t h = ctx->next_pk++;
    try {
        mg.team.run([&](int r) {
            set_device(mg.rk[r]);
            pk_upload_ranges(mg.rk[r].ctx, d, src, r < gp.active ? place[r] : none, h);
        });
    } catch (...) {
        for (int r = 0; r < mg.n; r++) {
            g16_ctx* c = mg.rk[r].ctx;
            auto it = c->pks.find(h);
            if (it != c->pks.end()) { cudaSetDevice(c->device); cudaDeviceSynchronize(); it->second.release(); c->pks.erase(it); }
        }
        cudaSetDevice(ctx->device);
        throw;
    }
    G16_CUDA(cudaSetDevice(ctx->device));
    mg.pks[h] = gp;
    *pk_handle = h;
}

void multi_pk_free(g16_ctx* ctx, uint64_t pk_handle) {
    MultiGpu& mg = *ctx->multi;
    if (!mg.pks.count(pk_handle)) throw std::runtime_error("unknown proving-key handle");
    for (int r = 0; r < mg.n; r++) {
        g16_ctx* c = mg.rk[r].ctx;
        G16_CUDA(cudaSetDevice(c->device));
        G16_CUDA(cudaDeviceSynchronize());
        auto it = c->pks.find(pk_handle);
        if (it != c->pks.end()) { it->second.release(); c->pks.erase(it); }
    }
    G16_CUDA(cudaSetDevice(ctx->device));
    mg.pks.erase(pk_handle);
}

void multi_pk_wire_range(g16_ctx* ctx, uint64_t pk_handle, uint64_t* lo, uint64_t* hi) {
    MultiGpu& mg = *ctx->multi;
    if (!mg.pks.count(pk_handle)) throw std::runtime_error("unknown proving-key handle");
    uint64_t l = ~0ull, h = 0;
    for (int r = 0; r < mg.n; r++) {
        const PkDev& pk = mg.rk[r].ctx->pks.at(pk_handle);
        if (pk.w_hi > pk.w_lo) { if (pk.w_lo < l) l = pk.w_lo; if (pk.w_hi > h) h = pk.w_hi; }
    }
    *lo = h > l ? l : 0; *hi = h > l ? h : 0;
}

const PkDev& multi_pk_of_rank(g16_ctx* ctx, uint64_t pk_handle, int rank) {
    MultiGpu& mg = *ctx->multi;
    if (!mg.pks.count(pk_handle)) throw std::runtime_error("unknown proving-key handle");
    if (rank < 0 || rank >= mg.n) throw std::runtime_error("rank out of range");
    return mg.rk[rank].ctx->pks.at(pk_handle);
}

void multi_r1cs_upload(g16_ctx* ctx, const g16_r1cs_desc* d, uint64_t handle) {
    MultiGpu& mg = *ctx->multi;
    try {
        mg.team.run([&](int r) {
            if (r == 0) return;                 // rank 0 = ctx itself: uploaded (and validated) by g16_r1cs_upload
            set_device(mg.rk[r]);
            upload_r1cs_to(mg.rk[r].ctx, d, handle);
        });
    } catch (...) {
        multi_r1cs_free(ctx, handle);
        throw;
    }
    G16_CUDA(cudaSetDevice(ctx->device));
}

void multi_r1cs_free(g16_ctx* ctx, uint64_t handle) {
    MultiGpu& mg = *ctx->multi;
    for (int r = 1; r < mg.n; r++) {
        g16_ctx* c = mg.rk[r].ctx;
        auto it = c->r1cs.find(handle);
        if (it == c->r1cs.end()) continue;
        cudaSetDevice(c->device);
        cudaDeviceSynchronize();
        it->second.release();
        c->r1cs.erase(it);
    }
    cudaSetDevice(ctx->device);
}

// =================================================================== prove
namespace {

struct ProveJob {
    MultiGpu* mg;
    uint64_t handle;
    const MultiIn* in;
    int L, g, X, rb;
    uint64_t n, n_local;
    uint64_t launches0;
};

// exchange k of the call in flight: this rank's arrays at `src` go, element by element, to their owners in the other
// distribution (buffer `target` of every rank); returns once this rank's stream waits for everybody's stores
void dist_exchange(const ProveJob& J, int r, int k, const Fr::T* src, int target, int count, int to_block) {
    MultiGpu& mg = *J.mg;
    RankState& R = mg.rk[r];
    DistExchange a;
    for (int p = 0; p < 8; p++) a.dst[p] = p < mg.n ? mg.rk[p].buf[target].as<Fr::T>() : nullptr;
    a.src = src; a.count = count; a.n_local = J.n_local; a.g = J.g; a.L = J.L; a.rb = J.rb; a.rank = (uint32_t)r; a.to_block = to_block;
    ntt_dist_exchange(a, R.hs);
    G16_CUDA(cudaEventRecord(R.ev_x[k], R.hs));
    if (!mg.bar.wait()) throw std::runtime_error("another rank failed");        // every rank has recorded its event
    for (int p = 0; p < mg.n; p++) G16_CUDA(cudaStreamWaitEvent(R.hs, mg.rk[p].ev_x[k], 0));
}
NttBatch dist_batch(Fr::T* p, int count, uint64_t n_local) { return NttBatch{{p, p + (count > 1 ? n_local : 0), p + (count > 2 ? 2 * n_local : 0)}, count}; }

// a, b, c of this rank in D(rb) in buf[0]  ->  this rank's block of h (bit-reversed order) in buf[1]; on R.hs
void dist_compute_h(const ProveJob& J, int r) {
    RankState& R = J.mg->rk[r];
    cudaStream_t hs = R.hs;
    NttDomain& dom = R.ctx->domain(J.L);
    Fr::T* A = R.buf[0].as<Fr::T>();
    Fr::T* B = R.buf[1].as<Fr::T>();
    const int g = J.g, X = J.X, rb = J.rb;
    const uint64_t nl = J.n_local;
    auto mark = [&](int i) { G16_CUDA(cudaEventRecord(R.mark[i], hs)); };
    mark(0);
    ntt_dist_stages(dom, dist_batch(A, 3, nl), TW_INV, true, g, (uint32_t)r, true, X, hs, rb);          // 3 x iNTT, upper stages
    mark(1);
    dist_exchange(J, r, 0, A, 1, 3, 1);
    mark(2);
    ntt_dist_stages(dom, dist_batch(B, 3, nl), TW_INV, true, g, (uint32_t)r, false, X, hs);             //           lower stages
    ntt_dist_stages(dom, dist_batch(B, 3, nl), TW_COSET_FWD, false, g, (uint32_t)r, false, X, hs);      // 3 x coset NTT, lower stages
    mark(3);
    dist_exchange(J, r, 1, B, 0, 3, 0);
    mark(4);
    ntt_dist_stages(dom, dist_batch(A, 3, nl), TW_COSET_FWD, false, g, (uint32_t)r, true, X, hs, rb);   //               upper stages
    fr_pointwise_h(A, A + nl, A + 2 * nl, nl, dom.h_c1, dom.h_c2, hs);
    ntt_dist_stages(dom, dist_batch(A, 1, nl), TW_COSET_INV, true, g, (uint32_t)r, true, X, hs, rb);    // coset iNTT, upper stages
    mark(5);
    dist_exchange(J, r, 2, A, 1, 1, 1);
    mark(6);
    ntt_dist_stages(dom, dist_batch(B, 1, nl), TW_COSET_INV, true, g, (uint32_t)r, false, X, hs);       //            lower stages
    mark(7);
}

void rank_prove(const ProveJob& J, int r) {
    MultiGpu& mg = *J.mg;
    RankState& R = mg.rk[r];
    g16_ctx* c = R.ctx;
    const MultiIn& in = *J.in;
    set_device(R);
    PkDev& pk = c->pks.at(J.handle);
    cudaStream_t st = c->stream, hs = R.hs;
    const int g = J.g, rb = J.rb, X = J.X;
    const uint64_t n = J.n, n_local = J.n_local;
    memset(&R.st, 0, sizeof(R.st));

    // buffers exist before anybody stores into them
    R.buf[0].reserve((size_t)3 * n_local * sizeof(Fr::T));
    R.buf[1].reserve((size_t)3 * n_local * sizeof(Fr::T));
    c->stage[3].reserve((size_t)pk.n_wires * 32);
    const R1csDev* rdev = nullptr;
    if (in.r1cs) {
        auto ir = c->r1cs.find(in.r1cs);
        if (ir == c->r1cs.end()) throw std::runtime_error("unknown R1CS handle");
        rdev = &ir->second;
        if (rdev->n_wires != pk.n_wires || rdev->n_constraints > n) throw std::runtime_error("constraint system does not match the proving key");
    }
    if (!mg.bar.wait()) throw std::runtime_error("another rank failed");

    G16_CUDA(cudaEventRecord(c->ev[26], st));
    if (in.device) {       // the caller's arrays are ready once rank 0's stream reaches ev_in (recorded before the ranks started)
        G16_CUDA(cudaStreamWaitEvent(st, mg.rk[0].ev_in, 0));
        G16_CUDA(cudaStreamWaitEvent(hs, mg.rk[0].ev_in, 0));
    }
    // ---- W.  A rank's pieces of A, B, K read wires from all over the vector (the placement cuts every array on its own), so
    // the ranks together would pull several copies of W through the host's memory system (measured at N = 8: 3.8 - 5.6 ms
    // until W was resident).  Instead every rank uploads ONE n-th of W and fetches the other slices it needs from its peers
    // over NVLink: the host serves W exactly once, in n parallel streams.
    const uint64_t w_lo = rdev ? 0 : pk.w_lo, w_hi = rdev ? pk.n_wires : pk.w_hi;
    const int dev0 = mg.rk[0].ctx->device;
    c->stage[3].reserve((size_t)pk.n_wires * 32);
    {
        uint64_t s_lo, s_hi;
        slice(pk.n_wires, (uint32_t)r, (uint32_t)mg.n, s_lo, s_hi);
        char* dst = (char*)c->stage[3].p + (size_t)s_lo * 32;
        const char* srcp = (const char*)in.W + (size_t)s_lo * 32;
        const size_t bytes = (size_t)(s_hi - s_lo) * 32;
        if (bytes) {
            if (in.device) G16_CUDA(cudaMemcpyPeerAsync(dst, c->device, srcp, dev0, bytes, st));
            else { h2d_caller(c, dst, srcp, bytes, st, mg.n); R.st.h2d_bytes += bytes; }
        }
        G16_CUDA(cudaEventRecord(R.ev_w, st));
    }
    if (!mg.bar.wait()) throw std::runtime_error("another rank failed");      // every rank has recorded its slice event
    for (int p = 0; p < mg.n; p++) {
        if (p == r) continue;
        uint64_t s_lo, s_hi;
        slice(pk.n_wires, (uint32_t)p, (uint32_t)mg.n, s_lo, s_hi);
        const uint64_t lo = s_lo > w_lo ? s_lo : w_lo, hi = s_hi < w_hi ? s_hi : w_hi;
        if (hi <= lo) continue;
        G16_CUDA(cudaStreamWaitEvent(st, mg.rk[p].ev_w, 0));
        G16_CUDA(cudaMemcpyPeerAsync((char*)c->stage[3].p + (size_t)lo * 32, c->device, (const char*)mg.rk[p].ctx->stage[3].p + (size_t)lo * 32,
                                     mg.rk[p].ctx->device, (size_t)(hi - lo) * 32, st));
    }
    const Fr::T* W_ptr = c->stage[3].as<Fr::T>();
    G16_CUDA(cudaEventRecord(c->ev[15], st));

    // ---- a, b, c of this rank in D(rb)
    Fr::T* A = R.buf[0].as<Fr::T>();
    Fr::T* B = R.buf[1].as<Fr::T>();
    if (rdev) {
        G16_CUDA(cudaStreamWaitEvent(hs, c->ev[15], 0));
        r1cs_eval(*rdev, W_ptr, A, A + n_local, A + 2 * n_local, n, hs, rb, g, (uint32_t)r);
    } else if (in.device) {
        PullArgs pa;
        pa.src[0] = (const Fr::T*)in.a; pa.src[1] = (const Fr::T*)in.b; pa.src[2] = (const Fr::T*)in.c;
        pa.dst = A; pa.n_local = n_local; pa.count = 3; pa.rb = rb; pa.g = g; pa.rank = (uint32_t)r;
        unsigned blocks = (unsigned)((3 * n_local + 255) / 256);
        if (blocks > 148 * 16) blocks = 148 * 16;
        k_dist_pull<<<blocks, 256, 0, hs>>>(pa);
        G16_LAUNCH_CHECK();
    } else {
        const void* srcs[3] = {in.a, in.b, in.c};
        for (int i = 0; i < 3; i++) load_host_rows(c, mg.n, A + i * n_local, srcs[i], in.n_constraints, n, n_local, rb, g, (uint32_t)r, hs);
        uint64_t mine = 0;   // elements of one array that exist on this rank
        for (uint64_t hi = 0; hi < (n_local >> rb); hi++) {
            const uint64_t s = (hi << (rb + g)) + ((uint64_t)r << rb);
            if (s >= in.n_constraints) break;
            mine += in.n_constraints - s < (1ull << rb) ? in.n_constraints - s : (1ull << rb);
        }
        R.st.h2d_bytes += 3 * mine * 32;
    }
    G16_CUDA(cudaEventRecord(c->ev[27], hs));

    // ---- distributed computeH (prove.go:356-399): 3 exchanges
    G16_CUDA(cudaEventRecord(c->ev[30], hs));
    dist_compute_h(J, r);
    G16_CUDA(cudaEventRecord(R.ev_x[3], hs));   // this rank's block of h (bit-reversed order) is final in buf[1]
    G16_CUDA(cudaEventRecord(c->ev[31], hs));
    // ---- the MSMs over W, on their own streams.  Their sorts (short CTAs) run beside computeH; their accumulations --
    // CTAs that own an SM for milliseconds, which a high-priority stream cannot displace -- are held back until this
    // rank's share of computeH is through: the other GPUs wait for its exchanges (measured at N = 4: h arrived after
    // 15 ms instead of < 3 when the accumulations were let loose first)
    prove_device(c, pk, W_ptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, c->ev[15], true, 1, nullptr, R.ev_x[3]);
    if (!mg.bar.wait()) throw std::runtime_error("another rank failed");

    // ---- h under this rank's slice of G1.Z
    const Fr::T* h_ptr = B;
    if (pk.n_z) {
        const uint64_t z_lo = pk.z_off, z_hi = pk.z_off + pk.n_z;
        const uint64_t p0 = z_lo / n_local, p1 = (z_hi - 1) / n_local;
        if (p0 == p1 && (int)p0 == r) {
            G16_CUDA(cudaStreamWaitEvent(st, R.ev_x[3], 0));
            h_ptr = B + (z_lo - p0 * n_local);
        } else {
            R.hbuf.reserve((size_t)pk.n_z * sizeof(Fr::T));
            for (uint64_t p = p0; p <= p1; p++) {
                const uint64_t s = z_lo > p * n_local ? z_lo : p * n_local, e = z_hi < (p + 1) * n_local ? z_hi : (p + 1) * n_local;
                G16_CUDA(cudaStreamWaitEvent(st, mg.rk[p].ev_x[3], 0));
                G16_CUDA(cudaMemcpyPeerAsync(R.hbuf.as<Fr::T>() + (s - z_lo), c->device, mg.rk[p].buf[1].as<Fr::T>() + (s - p * n_local),
                                             mg.rk[p].ctx->device, (size_t)(e - s) * sizeof(Fr::T), st));
            }
            h_ptr = R.hbuf.as<Fr::T>();
        }
    }
    // ---- H.Z MSM, wait for all five, window sums -> XYZZ partial results
    prove_device(c, pk, W_ptr, const_cast<Fr::T*>(h_ptr), nullptr, nullptr, nullptr, &R.st, J.launches0, nullptr, true, 2, &R.part);
    G16_CUDA(cudaStreamSynchronize(hs));
    G16_CUDA(cudaEventElapsedTime(&R.st.ms_compute_h, c->ev[30], c->ev[31]));
    G16_CUDA(cudaEventElapsedTime(&R.st.ms_total, c->ev[26], c->ev[14]));
    static const bool timeline = getenv("G16_TIMELINE") != nullptr;   // diagnostic: where this rank's share of computeH goes
    if (timeline) {
        float m[8], w = 0, e0 = 0;
        for (int i = 0; i < 8; i++) cudaEventElapsedTime(&m[i], c->ev[26], R.mark[i]);
        cudaEventElapsedTime(&w, c->ev[26], c->ev[15]);
        cudaEventElapsedTime(&e0, c->ev[26], c->ev[0]);
        fprintf(stderr, "[g16 timeline] rank %d: W ready %.2f | computeH: load done %.2f, iNTT-upper %.2f, xchg0 %.2f, lower+coset-lower %.2f, xchg1 %.2f, "
                        "upper+pointwise+upper %.2f, xchg2 %.2f, lower %.2f | MSM enqueue at %.2f | end %.2f\n",
                r, w, m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], e0, R.st.ms_total);
    }
}

}  // namespace

void multi_prove(g16_ctx* ctx, uint64_t pk_handle, const MultiIn& in, void* out_points, g16_stats* stats) {
    MultiGpu& mg = *ctx->multi;
    auto ig = mg.pks.find(pk_handle);
    if (ig == mg.pks.end()) throw std::runtime_error("unknown proving-key handle");
    const GroupPk& gp = ig->second;
    if (!in.device && in.n_wires != gp.n_wires) throw std::runtime_error("witness length does not match the proving key");
    if (!in.device && !in.r1cs && in.n_constraints > gp.domain_n) throw std::runtime_error("more constraints than the domain holds");
    if (gp.active == 1) {
        // a domain too small to distribute: the whole key lives on rank 0, which proves alone
        ctx->multi = nullptr;
        int rc;
        if (in.r1cs) rc = g16_prove_bn254_wires(ctx, pk_handle, in.r1cs, in.W, in.n_wires, out_points, stats);
        else if (in.device) rc = g16_prove_bn254_dev(ctx, pk_handle, in.W, (void*)in.a, (void*)in.b, (void*)in.c, out_points, stats);
        else rc = g16_prove_bn254(ctx, pk_handle, in.W, in.n_wires, in.a, in.b, in.c, in.n_constraints, out_points, stats);
        ctx->multi = &mg;
        if (rc) throw std::runtime_error(ctx->err);
        return;
    }
    ProveJob J;
    J.mg = &mg; J.handle = pk_handle; J.in = &in;
    J.n = gp.domain_n; J.L = ilog2_ceil(J.n); J.g = mg.g; J.X = split_stage(J.L, J.g); J.rb = J.X - J.g;
    J.n_local = J.n >> J.g;
    J.launches0 = launch_counter().load();
    if (in.device) G16_CUDA(cudaEventRecord(mg.rk[0].ev_in, ctx->stream));
    mg.bar.reset(mg.n);
    const auto t0 = std::chrono::steady_clock::now();
    try {
        mg.team.run([&](int r) {
            try { rank_prove(J, r); }
            catch (...) { mg.bar.fail(); cudaDeviceSynchronize(); cudaGetLastError(); throw; }
        });
    } catch (...) {
        cudaSetDevice(ctx->device);
        throw;
    }
    G16_CUDA(cudaSetDevice(ctx->device));
    // sum the ranks' partial results in XYZZ form, one inversion per point
    const auto t1 = std::chrono::steady_clock::now();
    char* out = (char*)out_points;
    for (int i = 0; i < 4; i++) {
        G1::Xyzz acc = mg.rk[0].part.g1[i];
        for (int r = 1; r < mg.n; r++) G1::add(acc, mg.rk[r].part.g1[i]);
        copy_out(out + G1SZ * i, G1::to_aff(acc));
    }
    {
        G2::Xyzz acc = mg.rk[0].part.g2;
        for (int r = 1; r < mg.n; r++) G2::add(acc, mg.rk[r].part.g2);
        copy_out(out + 4 * G1SZ, G2::to_aff(acc));
    }
    const auto t2 = std::chrono::steady_clock::now();
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        for (int r = 0; r < mg.n; r++) {
            const g16_stats& s = mg.rk[r].st;
            if (s.ms_total > stats->ms_total) stats->ms_total = s.ms_total;
            if (s.ms_compute_h > stats->ms_compute_h) stats->ms_compute_h = s.ms_compute_h;
            if (s.ms_host > stats->ms_host) stats->ms_host = s.ms_host;
            for (int i = 0; i < 5; i++) {
                if (s.ms_msm[i] > stats->ms_msm[i]) stats->ms_msm[i] = s.ms_msm[i];
                if (s.ms_accumulate[i] > stats->ms_accumulate[i]) stats->ms_accumulate[i] = s.ms_accumulate[i];
                if (s.msm_n[i] && (!stats->msm_c[i] || s.msm_c[i] < stats->msm_c[i])) { stats->msm_c[i] = s.msm_c[i]; stats->msm_windows[i] = s.msm_windows[i]; }
                stats->msm_n[i] += s.msm_n[i];
            }
            stats->h2d_bytes += s.h2d_bytes;
            stats->d2h_bytes += s.d2h_bytes;
        }
        stats->ms_host += std::chrono::duration<float, std::milli>(t2 - t1).count();
        stats->kernel_launches = (uint32_t)(launch_counter().load() - J.launches0);
        stats->ms_wall = std::chrono::duration<float, std::milli>(t2 - t0).count();   // wall clock of the group call
    }
}

// =================================================================== group primitives (BASELINE configs 2 and 3 on N GPUs)
namespace {

void launch_pull(const void* a, const void* b, const void* c, Fr::T* dst, int count, const ProveJob& J, int r, cudaStream_t st) {
    PullArgs pa;
    pa.src[0] = (const Fr::T*)a; pa.src[1] = (const Fr::T*)b; pa.src[2] = (const Fr::T*)c;
    pa.dst = dst; pa.n_local = J.n_local; pa.count = count; pa.rb = J.rb; pa.g = J.g; pa.rank = (uint32_t)r;
    unsigned blocks = (unsigned)(((uint64_t)count * J.n_local + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_dist_pull<<<blocks, 256, 0, st>>>(pa);
    G16_LAUNCH_CHECK();
}

// dst[i(q)] = src[q]: this rank's D(rb) positions back into a whole array on another (or the same) GPU
__global__ void __launch_bounds__(256) k_dist_push(const Fr::T* __restrict__ src, Fr::T* __restrict__ dst, uint64_t n_local, int rb, int g, uint32_t rank) {
    const uint64_t rmask = (1ull << rb) - 1;
    for (uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_local; q += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = ((q >> rb) << (rb + g)) | ((uint64_t)rank << rb) | (q & rmask);
        const uint4* s = reinterpret_cast<const uint4*>(src + q);
        uint4* d = reinterpret_cast<uint4*>(dst + i);
        const uint4 lo = s[0], hi = s[1];
        d[0] = lo; d[1] = hi;
    }
}

ProveJob geometry(MultiGpu& mg, uint64_t n) {
    ProveJob J;
    J.mg = &mg; J.handle = 0; J.in = nullptr; J.launches0 = 0;
    J.n = n; J.L = ilog2_ceil(n); J.g = mg.g; J.X = split_stage(J.L, J.g); J.rb = J.X - J.g;
    J.n_local = n >> mg.g;
    return J;
}

template <class F>
void run_ranks(g16_ctx* ctx, F&& body) {
    MultiGpu& mg = *ctx->multi;
    mg.bar.reset(mg.n);
    try {
        mg.team.run([&](int r) {
            try { set_device(mg.rk[r]); body(r); }
            catch (...) { mg.bar.fail(); cudaDeviceSynchronize(); cudaGetLastError(); throw; }
        });
    } catch (...) {
        cudaSetDevice(ctx->device);
        throw;
    }
    G16_CUDA(cudaSetDevice(ctx->device));
}

}  // namespace

bool multi_can_distribute(g16_ctx* ctx, uint64_t n) { return ctx->multi && is_pow2(n) && ilog2_ceil(n) >= 2 * ctx->multi->g + 4; }

// computeH with a_d, b_d, c_d (n elements each) on rank 0's GPU; h (bit-reversed order) is written back into a_d
void multi_compute_h_dev(g16_ctx* ctx, void* a_d, void* b_d, void* c_d, uint64_t n) {
    MultiGpu& mg = *ctx->multi;
    ProveJob J = geometry(mg, n);
    G16_CUDA(cudaEventRecord(mg.rk[0].ev_in, ctx->stream));
    run_ranks(ctx, [&](int r) {
        RankState& R = mg.rk[r];
        R.buf[0].reserve((size_t)3 * J.n_local * sizeof(Fr::T));
        R.buf[1].reserve((size_t)3 * J.n_local * sizeof(Fr::T));
        R.ctx->domain(J.L).table(TW_INV, R.hs); R.ctx->domain(J.L).table(TW_COSET_FWD, R.hs); R.ctx->domain(J.L).table(TW_COSET_INV, R.hs);
        if (!mg.bar.wait()) throw std::runtime_error("another rank failed");
        G16_CUDA(cudaStreamWaitEvent(R.hs, mg.rk[0].ev_in, 0));
        launch_pull(a_d, b_d, c_d, R.buf[0].as<Fr::T>(), 3, J, r, R.hs);
        G16_CUDA(cudaEventRecord(R.ev_x[3], R.hs));
        if (!mg.bar.wait()) throw std::runtime_error("another rank failed");
        for (int p = 0; p < mg.n; p++) G16_CUDA(cudaStreamWaitEvent(R.hs, mg.rk[p].ev_x[3], 0));   // everybody has read a_d before h lands in it
        dist_compute_h(J, r);
        G16_CUDA(cudaMemcpyPeerAsync((Fr::T*)a_d + (uint64_t)r * J.n_local, mg.rk[0].ctx->device, R.buf[1].p, R.ctx->device, (size_t)J.n_local * sizeof(Fr::T), R.hs));
        G16_CUDA(cudaStreamSynchronize(R.hs));
    });
}

// one transform of fft.Domain on an array that lives on rank 0's GPU, in place: FFTInverse(DIF) and FFT(DIT), with or
// without the coset (the shapes computeH is made of); returns false for the shapes that are not distributed
bool multi_ntt_dev(g16_ctx* ctx, void* data_d, uint64_t n, int inverse, int coset, int decimation) {
    const bool inv_dif = inverse && decimation == G16_DIF, fwd_dit = !inverse && decimation == G16_DIT;
    if (!inv_dif && !fwd_dit) return false;
    MultiGpu& mg = *ctx->multi;
    ProveJob J = geometry(mg, n);
    const NttTable tab = inverse ? (coset ? TW_COSET_INV : TW_INV) : (coset ? TW_COSET_FWD : TW_FWD);
    G16_CUDA(cudaEventRecord(mg.rk[0].ev_in, ctx->stream));
    run_ranks(ctx, [&](int r) {
        RankState& R = mg.rk[r];
        cudaStream_t hs = R.hs;
        R.buf[0].reserve((size_t)3 * J.n_local * sizeof(Fr::T));
        R.buf[1].reserve((size_t)3 * J.n_local * sizeof(Fr::T));
        NttDomain& dom = R.ctx->domain(J.L);
        dom.table(tab, hs);
        if (!mg.bar.wait()) throw std::runtime_error("another rank failed");
        G16_CUDA(cudaStreamWaitEvent(hs, mg.rk[0].ev_in, 0));
        Fr::T* A = R.buf[0].as<Fr::T>();
        Fr::T* B = R.buf[1].as<Fr::T>();
        Fr::T* whole = (Fr::T*)data_d;
        const int dev0 = mg.rk[0].ctx->device;
        if (inv_dif) {
            launch_pull(whole, whole, whole, A, 1, J, r, hs);
            G16_CUDA(cudaEventRecord(R.ev_x[3], hs));
            if (!mg.bar.wait()) throw std::runtime_error("another rank failed");
            for (int p = 0; p < mg.n; p++) G16_CUDA(cudaStreamWaitEvent(hs, mg.rk[p].ev_x[3], 0));   // all reads of the array precede the write-back
            ntt_dist_stages(dom, dist_batch(A, 1, J.n_local), tab, true, J.g, (uint32_t)r, true, J.X, hs, J.rb);
            dist_exchange(J, r, 0, A, 1, 1, 1);
            ntt_dist_stages(dom, dist_batch(B, 1, J.n_local), tab, true, J.g, (uint32_t)r, false, J.X, hs);
            fr_scale_powers(B, J.n_local, J.L - J.g, dom.n_inv, Fr::one(), 0, 0, hs);
            G16_CUDA(cudaMemcpyPeerAsync(whole + (uint64_t)r * J.n_local, dev0, B, R.ctx->device, (size_t)J.n_local * sizeof(Fr::T), hs));
        } else {
            G16_CUDA(cudaMemcpyPeerAsync(B, R.ctx->device, whole + (uint64_t)r * J.n_local, dev0, (size_t)J.n_local * sizeof(Fr::T), hs));
            G16_CUDA(cudaEventRecord(R.ev_x[3], hs));
            if (!mg.bar.wait()) throw std::runtime_error("another rank failed");
            for (int p = 0; p < mg.n; p++) G16_CUDA(cudaStreamWaitEvent(hs, mg.rk[p].ev_x[3], 0));
            ntt_dist_stages(dom, dist_batch(B, 1, J.n_local), tab, false, J.g, (uint32_t)r, false, J.X, hs);
            dist_exchange(J, r, 0, B, 0, 1, 0);
            ntt_dist_stages(dom, dist_batch(A, 1, J.n_local), tab, false, J.g, (uint32_t)r, true, J.X, hs, J.rb);
            unsigned blocks = (unsigned)((J.n_local + 255) / 256);
            if (blocks > 148 * 16) blocks = 148 * 16;
            k_dist_push<<<blocks, 256, 0, hs>>>(A, whole, J.n_local, J.rb, J.g, (uint32_t)r);
            G16_LAUNCH_CHECK();
        }
        G16_CUDA(cudaStreamSynchronize(hs));
    });
    return true;
}

// resident bases sharded by point range over the ranks (SURVEY.md 8(e) level 2), each slice with its own table
void multi_bases_upload(g16_ctx* ctx, const void* bases, uint64_t n, int g2, int from_device, uint64_t* handle) {
    MultiGpu& mg = *ctx->multi;
    const uint64_t h = ctx->next_pk++;
    const size_t sz = g2 ? G2SZ : G1SZ;
    try {
        run_ranks(ctx, [&](int r) {
            uint64_t lo, hi;
            slice(n, (uint32_t)r, (uint32_t)mg.n, lo, hi);
            bases_upload_one(mg.rk[r].ctx, (const char*)bases + lo * sz, hi - lo, g2 != 0, from_device != 0, mg.rk[0].ctx->device, h);
        });
    } catch (...) {
        for (int r = 0; r < mg.n; r++) {
            g16_ctx* c = mg.rk[r].ctx;
            auto it = c->bases.find(h);
            if (it != c->bases.end()) { cudaSetDevice(c->device); cudaDeviceSynchronize(); it->second.table.release(); c->bases.erase(it); }
        }
        cudaSetDevice(ctx->device);
        throw;
    }
    GroupBases gb; gb.n = n; gb.g2 = g2 != 0;
    mg.bases[h] = gb;
    *handle = h;
}

bool multi_has_bases(g16_ctx* ctx, uint64_t handle) { return ctx->multi && ctx->multi->bases.count(handle); }

void multi_bases_free(g16_ctx* ctx, uint64_t handle) {
    MultiGpu& mg = *ctx->multi;
    for (int r = 0; r < mg.n; r++) {
        g16_ctx* c = mg.rk[r].ctx;
        auto it = c->bases.find(handle);
        if (it == c->bases.end()) continue;
        cudaSetDevice(c->device);
        cudaDeviceSynchronize();
        it->second.table.release();
        c->bases.erase(it);
    }
    cudaSetDevice(ctx->device);
    mg.bases.erase(handle);
}

// scalars: n Fr on rank 0's GPU (on_device) or in host memory; every rank runs the MSM over its slice, the host adds
void multi_msm_resident(g16_ctx* ctx, uint64_t handle, const void* scalars, bool on_device, void* out) {
    MultiGpu& mg = *ctx->multi;
    const GroupBases& gb = mg.bases.at(handle);
    if (gb.n && !scalars) throw std::runtime_error("null argument");
    if (on_device) G16_CUDA(cudaEventRecord(mg.rk[0].ev_in, ctx->stream));
    run_ranks(ctx, [&](int r) {
        RankState& R = mg.rk[r];
        g16_ctx* c = R.ctx;
        const g16_ctx::Bases& b = c->bases.at(handle);
        uint64_t lo, hi;
        slice(gb.n, (uint32_t)r, (uint32_t)mg.n, lo, hi);
        cudaStream_t st = c->stream;
        const Fr::T* s_ptr = nullptr;
        if (b.n) {
            const char* src = (const char*)scalars + lo * 32;
            if (on_device) G16_CUDA(cudaStreamWaitEvent(st, mg.rk[0].ev_in, 0));
            if (on_device && c->device == mg.rk[0].ctx->device) s_ptr = (const Fr::T*)src;
            else {
                c->stage[3].reserve((size_t)b.n * 32);
                if (on_device) G16_CUDA(cudaMemcpyPeerAsync(c->stage[3].p, c->device, src, mg.rk[0].ctx->device, (size_t)b.n * 32, st));
                else h2d_caller(c, c->stage[3].p, src, (size_t)b.n * 32, st, mg.n);
                s_ptr = c->stage[3].as<Fr::T>();
            }
        }
        const int pre_c = (c->msm_c == 0 && c->msm_precompute) ? b.pre_c : 0;
        MsmPlan pl;
        if (gb.g2) {
            if (b.n) { msm_g2_enqueue(c->ws[4], s_ptr, nullptr, b.table.as<G2::Aff>(), b.n, c->msm_c, c->msm_waves, st, pl, nullptr, nullptr, pre_c); G16_CUDA(cudaStreamSynchronize(st)); }
            R.part.g2 = b.n ? msm_g2_finish_xyzz(c->ws[4], pl) : G2::inf();
        } else {
            if (b.n) { msm_g1_enqueue(c->ws[0], s_ptr, nullptr, b.table.as<G1::Aff>(), b.n, c->msm_c, c->msm_waves, st, pl, nullptr, nullptr, pre_c); G16_CUDA(cudaStreamSynchronize(st)); }
            R.part.g1[0] = b.n ? msm_g1_finish_xyzz(c->ws[0], pl) : G1::inf();
        }
    });
    if (gb.g2) {
        G2::Xyzz acc = mg.rk[0].part.g2;
        for (int r = 1; r < mg.n; r++) G2::add(acc, mg.rk[r].part.g2);
        copy_out(out, G2::to_aff(acc));
    } else {
        G1::Xyzz acc = mg.rk[0].part.g1[0];
        for (int r = 1; r < mg.n; r++) G1::add(acc, mg.rk[r].part.g1[0]);
        copy_out(out, G1::to_aff(acc));
    }
}

}  // namespace g16
