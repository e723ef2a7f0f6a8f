/* libg16b200 -- C ABI of the B200-native Groth16 (BN254) proving hot path.
 *
 * Drop-in boundary for the device library that gnark's accelerated prover binds through cgo
 * (reference: backend/groth16/bn254/icicle/icicle.go, which calls github.com/ingonyama-zk/iciclegnark).
 * Each entry point names the reference interface it replaces (file:line under /root/reference).
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; g16_last_error(ctx) describes the failure.
 *     There is NO CPU fallback: without a working CUDA device every compute entry point fails.
 *   - memory images are gnark-crypto's: fr/fp.Element = 4 little-endian uint64 limbs in Montgomery form
 *     (R = 2^256); G1Affine = X||Y (64 B); G2Affine = X.A0||X.A1||Y.A0||Y.A1 (128 B); infinity = all zero.
 *   - pointers are HOST pointers unless the function name ends in _dev, whose data pointers are DEVICE
 *     pointers on the context's GPU (results are still written to host memory).
 *   - a context is bound to its GPU(s) and is thread-compatible: one call in flight per context (callers that prove
 *     concurrently on one key serialise their calls or use one context per thread).
 *   - scalars equal to zero and bases equal to infinity are legal inputs everywhere; zero-length MSMs
 *     return infinity (the reference special-cases both on the host: icicle.go:98-105,330-333,343-347).
 */
#ifndef G16B200_H
#define G16B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct g16_ctx g16_ctx;

#define G16_CURVE_BN254 0
#define G16_CURVE_BLS12_377 1 /* accepted by g16_bls12_377_ctx_create only (g16b200_bls12_377.h) */

/* index of each MSM result in g16_prove_bn254's output block */
#define G16_OUT_AR 0      /* sum wA_j * G1.A_j            64 B at offset 0   (icicle.go:312-324 before the epilogue) */
#define G16_OUT_BS1 1     /* sum wB_j * G1.B_j            64 B at offset 64  (icicle.go:299-310) */
#define G16_OUT_KRS_K 2   /* sum w_priv * G1.K            64 B at offset 128 (icicle.go:337-362) */
#define G16_OUT_KRS_HZ 3  /* sum h_i * G1.Z_i             64 B at offset 192 (icicle.go:326-335) */
#define G16_OUT_BS2 4     /* sum wB_j * G2.B_j           128 B at offset 256 (icicle.go:377-393) */
#define G16_PROVE_OUT_BYTES 384

typedef struct {
    float ms_total;        /* device time of the whole call (CUDA events) */
    float ms_h2d;          /* host->device copies (host-pointer variants only) */
    float ms_compute_h;    /* the 7-transform quotient pipeline */
    float ms_msm[5];       /* indexed by G16_OUT_* */
    float ms_host;         /* host epilogue of the MSMs (Horner over windows + affine conversion), wall clock */
    int32_t msm_c[5];      /* window bits used */
    int32_t msm_windows[5];
    uint64_t msm_n[5];     /* points in each MSM (this shard) */
    uint64_t h2d_bytes;
    uint64_t d2h_bytes;
    uint32_t kernel_launches; /* kernels of this library launched by the call */
    uint32_t reserved;
    float ms_accumulate[5]; /* the bucket-accumulation kernel of each MSM (dominant kernel; roofline evidence) */
    float ms_wall;          /* multi-GPU contexts: host wall clock of the whole call (all ranks + the final sum) */
} g16_stats;

/* ---- context ------------------------------------------------------------------------------------- */
/* device_ids/n_devices: the GPUs this context drives, n_devices in {1, 2, 4, 8} (one context per process and GPU set;
 * replaces the implicit default-device state of iciclegnark -- icicle.go never passes a device id).  With more than
 * one device the context is a group: g16_pk_upload places slices of the key on every GPU (shard_count must be 0/1),
 * g16_prove_bn254 / _dev / _wires run the distributed computeH (exchanges = peer stores over NVLink, no NCCL) and every
 * GPU's share of the five MSMs inside the one call and return the COMBINED five points, exactly like a single-GPU
 * context; _dev pointers live on device_ids[0]; every other entry point (primitives, commitments) runs on
 * device_ids[0].  The devices must be peer-accessible.  The same ordinal may be listed more than once: the ranks then
 * share that GPU (functional emulation for tests; no speed-up). */
int g16_ctx_create(g16_ctx** out, int curve_id, const int* device_ids, int n_devices);
/* CUDA devices visible to this process (0 when there is none or the driver cannot be initialised) */
int g16_device_count(void);
void g16_ctx_destroy(g16_ctx* ctx);
const char* g16_last_error(const g16_ctx* ctx);
/* tunables: "msm_c" (0 = auto, 8..16), "msm_waves" (accumulate waves, default 4),
 * "streams" (1 = run the five MSMs back to back; default 5 = overlap the MSMs that do not depend on h),
 * "msm_precompute" (default 1: g16_pk_upload also keeps the 2^(c j) multiples of every key base resident -- the
 * analogue of icicle's precompute_factor -- so that all windows of a prove MSM share one bucket set and the window
 * can grow to 20 bits; costs 13x the key in HBM and is skipped when that exceeds half of the free memory; set 0
 * before the upload to skip the tables, or after it to run the classic per-window path on the same key),
 * "msm_precompute_g1" (default 1; 0 before the upload: only G2.B gets a table -- for keys whose tables and batch-affine
 * buffers do not fit the HBM together, e.g. BLS12-377 at 2^24: the G1 MSMs gain more from the batch-affine rounds),
 * "msm_pre_c" (window bits, 16..21, of the tables built by later uploads; 0 = chosen from the array size),
 * "stage_threads" (copier threads that stage PAGEABLE caller memory -- e.g. Go heap slices -- into the context's pinned
 * ring; 0 = half of the host cores divided by the GPU count, between 2 and 8; page-locked buffers are DMA'd directly),
 * "msm_ba_rounds" / "msm_ba_rounds_g1" / "msm_ba_rounds_g2" (rounds of batch-affine pair additions that pre-reduce the
 * sorted bucket entries before the XYZZ accumulation, default 4; 0 = off; applies to MSMs of at least 2^17 points; an upper
 * bound: small MSMs take fewer -- G1 below 12 M / 20 M / 32 M bucket entries 1 / 2 / 3, G2 below 6 M / 32 M 2 / 3),
 * "g2_weight" (multi-GPU key placement: cost of a G2 point in hundredths of a G1 point; 0 = by group size, 275 / 300 / 350 for 2 / 4 / 8 GPUs; before the upload) */
int g16_ctx_set(g16_ctx* ctx, const char* key, int64_t value);
/* total kernels launched by this library in this process (bench.py's gpu_launches evidence) */
uint64_t g16_kernel_launch_count(void);

/* ---- device-resident proving key ------------------------------------------------------------------ */
/* Replaces pk.setupDevicePointers (icicle.go:31-130) and deviceInfo (icicle/provingkey.go:10-23):
 * bases are uploaded once, kept in Montgomery affine form (no host-side conversion), infinity entries
 * of K are kept (not stripped), the coset/twiddle tables are generated on the device. */
typedef struct {
    uint64_t domain_n;             /* pk.Domain.Cardinality (power of two) */
    const void* g1_a; uint64_t n_a;   /* pk.G1.A  (len nbWires - NbInfinityA) */
    const void* g1_b; uint64_t n_b;   /* pk.G1.B  (len nbWires - NbInfinityB) */
    const void* g1_k; uint64_t n_k;   /* pk.G1.K  (private wires) */
    const void* g1_z; uint64_t n_z;   /* pk.G1.Z  (domain_n - 1, bit-reversed at setup, setup.go:265-267) */
    const void* g2_b; uint64_t n_b2;  /* pk.G2.B  (same length as g1_b) */
    const uint8_t* infinity_a;     /* pk.InfinityA, one byte per wire (setup.go:212-237) */
    const uint8_t* infinity_b;     /* pk.InfinityB */
    uint64_t n_wires;
    const uint32_t* k_index;       /* wire index feeding each K base (after filterHeap, prove.go:241-246);
                                      NULL = wires [n_wires - n_k, n_wires) in order */
    uint32_t shard_index;          /* this context keeps slice shard_index of shard_count of every base */
    uint32_t shard_count;          /*   array (point-range sharding across GPUs); 0/1 = whole key */
} g16_pk_desc;
int g16_pk_upload(g16_ctx* ctx, const g16_pk_desc* desc, uint64_t* pk_handle);
/* SURVEY.md 8(f) N2 -- load the key straight from a ProvingKey.WriteDump image (marshal.go:386-456).  The dump's
 * point slices (unsafe.WriteSlice: u64 length, then the raw memory of the []G1Affine / []G2Affine) are already the
 * device format, so they never need to become Go slices: the caller decodes the small header with gnark's own
 * decoder (domain, alpha/beta/delta, InfinityA/B -- marshal.go:476-509), notes the byte offset of each slice's
 * first element, and the library streams file -> pinned staging -> HBM (two 32 MiB buffers, pread overlapped with
 * the H2D copy), reading only this shard's ranges.  `key` is a g16_pk_desc whose five base pointers are ignored. */
typedef struct {
    g16_pk_desc key;               /* counts, domain_n, infinity_a/b, n_wires, k_index, shard_*; g1_a..g2_b unused */
    int fd;                        /* open file, readable with pread(2) */
    uint64_t off_g1_a, off_g1_b, off_g1_k, off_g1_z, off_g2_b;   /* byte offset of element 0 of each slice */
} g16_pk_file_desc;
int g16_pk_upload_from_file(g16_ctx* ctx, const g16_pk_file_desc* desc, uint64_t* pk_handle);
int g16_pk_free(g16_ctx* ctx, uint64_t pk_handle);
/* The wires [*lo, *hi) of the witness vector are the only ones this context's shard of the A, B and K bases reads (the
 * whole vector for an unsharded key).  g16_prove_bn254 copies just that range; callers that keep W on the device
 * (g16_prove_bn254_dev, g16_prove_msms_dev) need to fill only that range. */
int g16_pk_wire_range(g16_ctx* ctx, uint64_t pk_handle, uint64_t* lo, uint64_t* hi);

/* Multi-GPU contexts: where the key lives.  lo[i], hi[i] (i = G16_OUT_*) receive the element range of G1.A, G1.B, G1.K,
 * G1.Z, G2.B that the group's rank `rank` (0 .. n_devices-1) keeps resident; empty ranges have lo == hi.  On a
 * single-GPU context rank must be 0 and the ranges are the context's shard (shard_index / shard_count).  Mirrors what
 * setupDevicePointers would log about its uploads (icicle.go:31-130). */
int g16_pk_placement(g16_ctx* ctx, uint64_t pk_handle, int rank, uint64_t lo[5], uint64_t hi[5]);

/* ---- the proof hot path ---------------------------------------------------------------------------- */
/* Replaces the device section of icicle_bn254.Prove (icicle.go:204-410): computeH + the five MSMs.
 * W: full wire vector (n_wires Fr); a,b,c: solver outputs L.w, R.w, O.w (n_constraints Fr each).
 * out_points: G16_PROVE_OUT_BYTES bytes, the five raw MSM results in affine form (this shard's partial
 * sums when the key is sharded).  The O(1) epilogue (icicle.go:281-295,306-307,319-321,362-372,386-391)
 * stays with the caller, or use g16_assemble_proof_bn254. */
int g16_prove_bn254(g16_ctx* ctx, uint64_t pk_handle, const void* W, uint64_t n_wires, const void* a,
                    const void* b, const void* c, uint64_t n_constraints, void* out_points, g16_stats* stats);
/* device-pointer variant: W_d (n_wires), a_d/b_d/c_d (domain_n elements each, zero padded, CLOBBERED) */
int g16_prove_bn254_dev(g16_ctx* ctx, uint64_t pk_handle, const void* W_d, void* a_d, void* b_d, void* c_d,
                        void* out_points, g16_stats* stats);

/* ---- a, b, c on the device ---------------------------------------------------------- */
/* The reference's solver produces a = L.w, b = R.w, c = O.w on the CPU (constraint/bn254/solver.go:540-628) and the
 * prover uploads all three (icicle.go:219,478-480: 3 x n x 32 B per proof).  With the constraint matrices resident
 * (uploaded once, like the key) only the wire vector crosses PCIe.  Matrices are CSR over the constraints; every
 * term is (coefficient id, wire id) exactly like gnark's constraint.Term{CID, VID} with the table cs.Coefficients. */
typedef struct {
    uint64_t n_constraints;
    uint64_t n_wires;
    const void* coeffs; uint64_t n_coeffs;     /* cs.Coefficients: Fr, Montgomery */
    const uint32_t* row_ptr[3];                /* L, R, O: n_constraints + 1 offsets into wire / coeff_id */
    const uint32_t* wire[3];                   /* Term.VID */
    const uint32_t* coeff_id[3];               /* Term.CID */
} g16_r1cs_desc;
int g16_r1cs_upload(g16_ctx* ctx, const g16_r1cs_desc* desc, uint64_t* r1cs_handle);
int g16_r1cs_free(g16_ctx* ctx, uint64_t r1cs_handle);
/* a_d, b_d, c_d: domain_n Fr each; rows >= n_constraints are zeroed (the padding of icicle.go:463-467) */
int g16_r1cs_eval_dev(g16_ctx* ctx, uint64_t r1cs_handle, const void* W_d, void* a_d, void* b_d, void* c_d,
                      uint64_t domain_n);
/* g16_prove_bn254 without a, b, c: uploads W (host, n_wires Fr), evaluates the matrices, proves */
int g16_prove_bn254_wires(g16_ctx* ctx, uint64_t pk_handle, uint64_t r1cs_handle, const void* W, uint64_t n_wires,
                          void* out_points, g16_stats* stats);

/* Host epilogue of prove.go:180-195,209-224,251-279,297-302 with injected randomness r, s (Fr, Montgomery):
 * pk_points = G1.Alpha || G1.Beta || G1.Delta (3 x 64 B) || G2.Beta || G2.Delta (2 x 128 B);
 * proof_out = Ar (64) || Bs (128) || Krs (64). */
int g16_assemble_proof_bn254(const void* pk_points, const void* r, const void* s, const void* msm_points,
                             void* proof_out);

/* ---- primitives (BASELINE configs 2 and 3, parity tests) ------------------------------------------ */
/* replaces iciclegnark.MsmOnDevice (icicle.go:302,315,332,355) / curve.G1Jac.MultiExp (prove.go:204-247) */
int g16_msm_g1(g16_ctx* ctx, const void* scalars, const void* bases, uint64_t n, void* out_affine);
int g16_msm_g1_dev(g16_ctx* ctx, const void* scalars_d, const void* bases_d, uint64_t n, void* out_affine);
/* Resident bases (the Groth16 case: bases are key material, fixed across proofs; the analogue of icicle's
 * precompute_bases + MSMConfig.precompute_factor): upload once, optionally with the table of 2^(c j) multiples, then
 * run MSMs against the handle with device-resident scalars.  g2 != 0: G2 bases / result. */
int g16_bases_upload(g16_ctx* ctx, const void* bases, uint64_t n, int g2, int from_device, uint64_t* bases_handle);
int g16_bases_free(g16_ctx* ctx, uint64_t bases_handle);
int g16_msm_resident_dev(g16_ctx* ctx, uint64_t bases_handle, const void* scalars_d, void* out_affine);
int g16_msm_resident(g16_ctx* ctx, uint64_t bases_handle, const void* scalars, void* out_affine);   /* host scalars */
/* BSB22 commitment keys: pedersen.ProvingKey{Basis, BasisExpSigma} (setup.go:57,295-305; dump
 * layout marshal.go:447-452) resident on the device, n G1 affine points each ((0,0) = infinity is legal).
 * g16_commit replaces pk.CommitmentKeys[i].Commit(values) inside the bsb22 solver hint (prove.go:94; icicle.go keeps
 * the CPU call) and pk.CommitmentKeys[i].ProveKnowledge(values) (prove.go:124): commitment_out = sum values_j Basis_j,
 * pok_out = sum values_j BasisExpSigma_j, both Montgomery affine; either may be NULL.  values: n_values Fr,
 * Montgomery, host memory; n_values must equal the key's length ("must have as many values as basis elements").
 * An empty key yields the point at infinity.  Hashing the commitment to the wire value and folding the proofs of
 * knowledge (prove.go:97-108,125-135) are O(#commitments) host work and stay with the caller. */
int g16_commitment_key_upload(g16_ctx* ctx, const void* basis, const void* basis_exp_sigma, uint64_t n,
                              uint64_t* key_handle);
int g16_commitment_key_free(g16_ctx* ctx, uint64_t key_handle);
int g16_commit(g16_ctx* ctx, uint64_t key_handle, const void* values, uint64_t n_values, void* commitment_out,
               void* pok_out);
/* replaces iciclegnark.MsmG2OnDevice (icicle.go:382) / curve.G2Jac.MultiExp (prove.go:293) */
int g16_msm_g2(g16_ctx* ctx, const void* scalars, const void* bases, uint64_t n, void* out_affine);
int g16_msm_g2_dev(g16_ctx* ctx, const void* scalars_d, const void* bases_d, uint64_t n, void* out_affine);

/* fft.Domain transforms on n = 2^k Fr elements, in place (prove.go:372-378,396; icicle.go:489,490,502,510).
 * inverse: 0 = FFT, 1 = FFTInverse (scaled by 1/n);  coset: 1 = fft.OnCoset() (shift g = 5);
 * decimation: G16_DIF = natural in -> bit-reversed out, G16_DIT = bit-reversed in -> natural out. */
#define G16_DIF 0
#define G16_DIT 1
int g16_ntt(g16_ctx* ctx, void* data, uint64_t n, int inverse, int coset, int decimation);
int g16_ntt_dev(g16_ctx* ctx, void* data_d, uint64_t n, int inverse, int coset, int decimation);

/* computeH (prove.go:356-399 / icicle.go:453-513): h_out receives domain_n Fr in bit-reversed order */
int g16_compute_h(g16_ctx* ctx, const void* a, const void* b, const void* c, uint64_t n_constraints,
                  uint64_t domain_n, void* h_out);
/* a_d,b_d,c_d: domain_n elements each (zero padded); h is left in a_d, b_d/c_d are clobbered */
int g16_compute_h_dev(g16_ctx* ctx, void* a_d, void* b_d, void* c_d, uint64_t domain_n);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  computeH like
 * g16_compute_h_dev, but every byte of scratch (the zero-padded working copies of a, b, c and the three twiddle
 * tables) lives in a static device array inside the library: 6 x G16_NOALLOC_MAX_DOMAIN x 32 B = 12 MiB, held by
 * every GPU that loads the library until the process exits.  The call performs no cudaMalloc, cudaFree or other
 * allocating routine.  a_d, b_d, c_d: n_constraints Fr each on the context's GPU, left untouched (no padding needed);
 * h_out_d: domain_n Fr, bit-reversed order, must not overlap the inputs.  domain_n > G16_NOALLOC_MAX_DOMAIN is
 * rejected with an error.  Every call on a GPU shares the one array: calls that overlap on the GPU (two contexts,
 * two threads) race -- serialise them.  Returns after the context's stream has drained. */
#define G16_NOALLOC_MAX_DOMAIN 65536
int g16_compute_h_noalloc_dev(g16_ctx* ctx, const void* a_d, const void* b_d, const void* c_d, uint64_t n_constraints,
                              uint64_t domain_n, void* h_out_d);

/* The three g16_dist_*_dev calls below return after ctx's stream has drained, unless g16_ctx_set(ctx, "dist_sync", 0):
 * then they only enqueue, and the caller orders its exchanges against g16_ctx_stream(ctx) (e.g. by issuing the NCCL
 * all-to-all on that same stream). */
/* ---- distributed computeH (one context per GPU; the caller moves data between ranks, e.g. NCCL all-to-all) ----
 * A size-n transform over world = 2^g ranks is split at global stage `split` (g <= split <= log2(n) - g): stages
 * with half-size >= 2^split run on the CYCLIC distribution (rank r holds positions q*world + r), the others on
 * the BLOCK distribution (rank r holds [r n/world, (r+1) n/world)); one all-to-all in between -- the four-step
 * transpose of SURVEY.md 8(e).  data_d = `count` arrays of n/world elements, back to back.
 * table: 0 forward, 1 inverse, 2 coset forward, 3 coset inverse (no 1/n scaling: computeH folds it) */
int g16_dist_ntt_stages_dev(g16_ctx* ctx, void* data_d, int count, uint64_t n_global, int world, int rank, int table,
                            int dif, int cyclic, int split);
/* layout change around the all-to-all.  Wire layout = [peer][array][q'], q' < n_local/world;
 * block_layout: 0 = local arrays are cyclic-distributed, 1 = block-distributed; to_wire: 1 pack, 0 unpack */
int g16_dist_permute_dev(g16_ctx* ctx, const void* src_d, void* dst_d, int count, uint64_t n_local, int world,
                         int block_layout, int to_wire);
/* (a*b/n - c) * den/n^2 on this rank's elements; abc_d = a, b, c back to back, result in a (icicle.go:500) */
int g16_dist_pointwise_h_dev(g16_ctx* ctx, void* abc_d, uint64_t n_local, uint64_t n_global);
/* the five MSMs only; h_local_d = this shard's block of h (domain_n/shard_count elements, bit-reversed order) */
int g16_prove_msms_dev(g16_ctx* ctx, uint64_t pk_handle, const void* W_d, const void* h_local_d, void* out_points,
                       g16_stats* stats);
/* Optional first half of g16_prove_msms_dev: enqueues the four MSMs over W (ar, bs1, krs_k, bs2) on their own streams and
 * returns at once, so that they run while the caller produces h (the distributed computeH and its all-to-alls).  The next
 * g16_prove_msms_dev on the same key then only adds the H.Z MSM and collects all five.  W_d must stay valid until then. */
int g16_prove_msms_begin_dev(g16_ctx* ctx, uint64_t pk_handle, const void* W_d);

/* out[i] = a[i] * b[i] in Fr (icicle.go:500 PolyOps building block; minimum parity slice) */
int g16_fr_mul_vec(g16_ctx* ctx, const void* a, const void* b, uint64_t n, void* out);

/* ---- setup-side tooling ---------------------------------------------------------------------------- */
/* out[i] = [k_i] G with G the BN254 G1 / G2 generator: device analogue of curve.BatchScalarMultiplicationG1/G2
 * (setup.go:251,320), used to materialise proving keys at BASELINE sizes. */
int g16_g1_fixed_base_mul(g16_ctx* ctx, const void* scalars, uint64_t n, void* out_points);
int g16_g2_fixed_base_mul(g16_ctx* ctx, const void* scalars, uint64_t n, void* out_points);
int g16_g1_fixed_base_mul_dev(g16_ctx* ctx, const void* scalars_d, uint64_t n, void* out_points_d);
int g16_g2_fixed_base_mul_dev(g16_ctx* ctx, const void* scalars_d, uint64_t n, void* out_points_d);

/* host-side affine group helpers (epilogue arithmetic for non-Go hosts): out = a + b ; out = [k] a */
int g16_g1_add(const void* a, const void* b, void* out);
int g16_g2_add(const void* a, const void* b, void* out);
int g16_g1_scalar_mul(const void* a, const void* k_fr_mont, void* out);
int g16_g2_scalar_mul(const void* a, const void* k_fr_mont, void* out);

/* ---- measurement ----------------------------------------------------------------------------------- */
/* Measured integer-pipe peak on this GPU: independent 32-bit IMAD (mad.lo.u32) and IMAD.WIDE (mad.wide.u32)
 * issue rates in operations per second over all SMs -- the MSM/NTT roofline denominator. */
int g16_imad_peak(g16_ctx* ctx, double* imad_per_s, double* imad_wide_per_s);
/* out3[0] = IMAD.WIDE.U32.X (carry-chained, the form the Montgomery multiplier issues) rate per second;
 * out3[1], out3[2] = in-register Fr / Fp Montgomery multiplications per second (the practical multiplier peaks;
 * the two differ only for BLS12-377, whose Fp has 12 limbs). */
int g16_pipe_rates(g16_ctx* ctx, double* out3);
/* the library's stream, so callers can bracket calls with CUDA events */
void* g16_ctx_stream(g16_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif
