"""ctypes binding of libg16b200.so (the C ABI in include/g16b200.h).

This is the Python stand-in for the cgo layer a Go host would use (INTEGRATION.md shows the cgo stub);
it exists because neither build machine has a Go toolchain.  Arrays are numpy ``uint64`` with a trailing
limb axis -- Fr ``(n, 4)``, G1 affine ``(n, 8)``, G2 affine ``(n, 16)`` for BN254 (``(n, 12)`` / ``(n, 24)`` for
BLS12-377, ``Context(curve="bls12_377")``, entry points ``g16_bls12_377_*``) -- i.e. gnark-crypto's memory image.
Device-resident operands are passed as integer device pointers (``torch.Tensor.data_ptr()``); torch is
plumbing only.  There is no CPU fallback: a missing library or GPU raises.
"""
import ctypes
import os

import numpy as np

# a prove runs on 10 CUDA streams; the default 8 hardware queues would make two of them share one (see g16_ctx_create).
# Must be in the environment before the process initialises CUDA (torch does so lazily, after this import).
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libg16b200.so")

DIF, DIT = 0, 1
OUT_AR, OUT_BS1, OUT_KRS_K, OUT_KRS_HZ, OUT_BS2 = range(5)
PROVE_OUT_BYTES = 384

# curve -> (ctx_create curve_id, G1 affine u64 words, G2 affine u64 words)
CURVES = {"bn254": (0, 8, 16), "bls12_377": (1, 12, 24)}


class G16Error(RuntimeError):
    pass


class Stats(ctypes.Structure):
    _fields_ = [
        ("ms_total", ctypes.c_float), ("ms_h2d", ctypes.c_float), ("ms_compute_h", ctypes.c_float),
        ("ms_msm", ctypes.c_float * 5), ("ms_host", ctypes.c_float),
        ("msm_c", ctypes.c_int32 * 5), ("msm_windows", ctypes.c_int32 * 5), ("msm_n", ctypes.c_uint64 * 5),
        ("h2d_bytes", ctypes.c_uint64), ("d2h_bytes", ctypes.c_uint64),
        ("kernel_launches", ctypes.c_uint32), ("reserved", ctypes.c_uint32),
        ("ms_accumulate", ctypes.c_float * 5), ("ms_wall", ctypes.c_float),
    ]

    def as_dict(self):
        return dict(ms_total=self.ms_total, ms_h2d=self.ms_h2d, ms_compute_h=self.ms_compute_h,
                    ms_msm=list(self.ms_msm), ms_host=self.ms_host, msm_c=list(self.msm_c),
                    msm_windows=list(self.msm_windows), msm_n=list(self.msm_n), h2d_bytes=self.h2d_bytes,
                    d2h_bytes=self.d2h_bytes, kernel_launches=self.kernel_launches,
                    ms_accumulate=list(self.ms_accumulate), ms_wall=self.ms_wall)


class PkDesc(ctypes.Structure):
    _fields_ = [
        ("domain_n", ctypes.c_uint64),
        ("g1_a", ctypes.c_void_p), ("n_a", ctypes.c_uint64),
        ("g1_b", ctypes.c_void_p), ("n_b", ctypes.c_uint64),
        ("g1_k", ctypes.c_void_p), ("n_k", ctypes.c_uint64),
        ("g1_z", ctypes.c_void_p), ("n_z", ctypes.c_uint64),
        ("g2_b", ctypes.c_void_p), ("n_b2", ctypes.c_uint64),
        ("infinity_a", ctypes.c_void_p), ("infinity_b", ctypes.c_void_p),
        ("n_wires", ctypes.c_uint64),
        ("k_index", ctypes.c_void_p),
        ("shard_index", ctypes.c_uint32), ("shard_count", ctypes.c_uint32),
    ]


class PkFileDesc(ctypes.Structure):
    _fields_ = [
        ("key", PkDesc), ("fd", ctypes.c_int),
        ("off_g1_a", ctypes.c_uint64), ("off_g1_b", ctypes.c_uint64), ("off_g1_k", ctypes.c_uint64),
        ("off_g1_z", ctypes.c_uint64), ("off_g2_b", ctypes.c_uint64),
    ]


class R1csDesc(ctypes.Structure):
    _fields_ = [
        ("n_constraints", ctypes.c_uint64), ("n_wires", ctypes.c_uint64),
        ("coeffs", ctypes.c_void_p), ("n_coeffs", ctypes.c_uint64),
        ("row_ptr", ctypes.c_void_p * 3), ("wire", ctypes.c_void_p * 3), ("coeff_id", ctypes.c_void_p * 3),
    ]


# every symbol include/g16b200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "g16_ctx_create", "g16_device_count", "g16_ctx_destroy", "g16_last_error", "g16_ctx_set", "g16_kernel_launch_count",
    "g16_pk_upload", "g16_pk_upload_from_file", "g16_pk_free", "g16_pk_wire_range", "g16_pk_placement", "g16_prove_bn254", "g16_prove_bn254_dev", "g16_assemble_proof_bn254",
    "g16_r1cs_upload", "g16_r1cs_free", "g16_r1cs_eval_dev", "g16_prove_bn254_wires",
    "g16_bases_upload", "g16_bases_free", "g16_msm_resident_dev", "g16_msm_resident",
    "g16_commitment_key_upload", "g16_commitment_key_free", "g16_commit",
    "g16_msm_g1", "g16_msm_g1_dev", "g16_msm_g2", "g16_msm_g2_dev", "g16_ntt", "g16_ntt_dev",
    "g16_compute_h", "g16_compute_h_dev", "g16_compute_h_noalloc_dev", "g16_fr_mul_vec",
    "g16_dist_ntt_stages_dev", "g16_dist_permute_dev", "g16_dist_pointwise_h_dev", "g16_prove_msms_dev", "g16_prove_msms_begin_dev",
    "g16_g1_fixed_base_mul", "g16_g2_fixed_base_mul", "g16_g1_fixed_base_mul_dev", "g16_g2_fixed_base_mul_dev",
    "g16_g1_add", "g16_g2_add", "g16_g1_scalar_mul", "g16_g2_scalar_mul", "g16_imad_peak", "g16_pipe_rates", "g16_ctx_stream",
]

def symbol_name(name, curve="bn254"):
    """public name of entry point `name` (a BN254 name from SYMBOLS) for `curve` (tools/gen_curve_abi.py's rule)."""
    if curve == "bn254":
        return name
    assert curve == "bls12_377", curve
    return "g16_bls12_377_" + name[len("g16_"):].replace("_bn254", "")


class _CurveLib:
    """attribute access by BN254 name, resolved to the curve's own entry point."""

    def __init__(self, cdll, curve):
        self._cdll, self._curve = cdll, curve

    def __getattr__(self, name):
        fn = getattr(self._cdll, symbol_name(name, self._curve))
        setattr(self, name, fn)
        return fn


_libs = {}


def load(curve="bn254"):
    """dlopen the in-tree library; raises if it has not been built (no fallback)."""
    if curve not in CURVES:
        raise G16Error(f"unknown curve {curve!r} (have {sorted(CURVES)})")
    if curve not in _libs:
        if not os.path.exists(LIB_PATH):
            raise G16Error(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(make -C gnark_icicle_b200/csrc); there is no CPU fallback")
        lib = _CurveLib(ctypes.CDLL(LIB_PATH), curve)
        lib.g16_last_error.restype = ctypes.c_char_p
        lib.g16_last_error.argtypes = [ctypes.c_void_p]
        lib.g16_kernel_launch_count.restype = ctypes.c_uint64
        lib.g16_ctx_stream.restype = ctypes.c_void_p
        lib.g16_ctx_stream.argtypes = [ctypes.c_void_p]
        lib.g16_ctx_destroy.argtypes = [ctypes.c_void_p]
        lib.g16_ctx_destroy.restype = None
        _libs[curve] = lib
    return _libs[curve]


def _ptr(a):
    if a is None:
        return ctypes.c_void_p(0)
    if isinstance(a, int):
        return ctypes.c_void_p(a)
    return ctypes.c_void_p(a.ctypes.data)


def _arr(a, width):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if a.ndim == 1 and a.size == width:
        a = a.reshape(1, width)
    assert a.ndim == 2 and a.shape[1] == width, (a.shape, width)
    return a


u64 = ctypes.c_uint64


class Context:
    """One GPU -- or a group of 2 / 4 / 8 -- and one call in flight (g16_ctx)."""

    def __init__(self, device=0, curve="bn254"):
        """device: one ordinal, or a list of 2 / 4 / 8 ordinals = a multi-GPU group behind the same calls."""
        self._lib = load(curve)
        self.curve = curve
        curve_id, self.g1w, self.g2w = CURVES[curve]
        self.prove_out_words = 4 * self.g1w + self.g2w
        self._h = ctypes.c_void_p()
        devices = [int(device)] if isinstance(device, (int, np.integer)) else [int(d) for d in device]
        ids = (ctypes.c_int * len(devices))(*devices)
        rc = self._lib.g16_ctx_create(ctypes.byref(self._h), curve_id, ids, len(devices))
        if rc != 0:
            self._h = ctypes.c_void_p()
            raise G16Error(f"g16_ctx_create failed with status {rc} (no usable CUDA device, or devices that cannot map "
                           "each other? there is no CPU fallback)")
        self.device = devices[0]
        self.devices = devices

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.g16_ctx_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise G16Error(f"{what}: {self._lib.g16_last_error(self._h).decode()}")

    def set(self, key, value):
        self._check(self._lib.g16_ctx_set(self._h, key.encode(), ctypes.c_int64(value)), "g16_ctx_set")

    @property
    def stream(self):
        return self._lib.g16_ctx_stream(self._h)

    # ---- primitives, host arrays
    def fr_mul_vec(self, a, b):
        a, b = _arr(a, 4), _arr(b, 4)
        out = np.empty_like(a)
        self._check(self._lib.g16_fr_mul_vec(self._h, _ptr(a), _ptr(b), u64(len(a)), _ptr(out)), "g16_fr_mul_vec")
        return out

    def msm_g1(self, scalars, bases):
        s, p = _arr(scalars, 4), _arr(bases, self.g1w)
        assert len(s) == len(p)
        out = np.zeros(self.g1w, dtype=np.uint64)
        self._check(self._lib.g16_msm_g1(self._h, _ptr(s), _ptr(p), u64(len(s)), _ptr(out)), "g16_msm_g1")
        return out

    def msm_g2(self, scalars, bases):
        s, p = _arr(scalars, 4), _arr(bases, self.g2w)
        assert len(s) == len(p)
        out = np.zeros(self.g2w, dtype=np.uint64)
        self._check(self._lib.g16_msm_g2(self._h, _ptr(s), _ptr(p), u64(len(s)), _ptr(out)), "g16_msm_g2")
        return out

    def ntt(self, data, inverse=False, coset=False, decimation=DIF):
        x = _arr(data, 4).copy()
        self._check(self._lib.g16_ntt(self._h, _ptr(x), u64(len(x)), int(inverse), int(coset), int(decimation)), "g16_ntt")
        return x

    def compute_h(self, a, b, c, domain_n):
        a, b, c = _arr(a, 4), _arr(b, 4), _arr(c, 4)
        assert len(a) == len(b) == len(c)
        h = np.empty((domain_n, 4), dtype=np.uint64)
        self._check(self._lib.g16_compute_h(self._h, _ptr(a), _ptr(b), _ptr(c), u64(len(a)), u64(domain_n), _ptr(h)), "g16_compute_h")
        return h

    def g1_fixed_base_mul(self, scalars):
        s = _arr(scalars, 4)
        out = np.empty((len(s), self.g1w), dtype=np.uint64)
        self._check(self._lib.g16_g1_fixed_base_mul(self._h, _ptr(s), u64(len(s)), _ptr(out)), "g16_g1_fixed_base_mul")
        return out

    def g2_fixed_base_mul(self, scalars):
        s = _arr(scalars, 4)
        out = np.empty((len(s), self.g2w), dtype=np.uint64)
        self._check(self._lib.g16_g2_fixed_base_mul(self._h, _ptr(s), u64(len(s)), _ptr(out)), "g16_g2_fixed_base_mul")
        return out

    # ---- primitives, device pointers (ints); results on host
    def msm_g1_dev(self, scalars_ptr, bases_ptr, n):
        out = np.zeros(self.g1w, dtype=np.uint64)
        self._check(self._lib.g16_msm_g1_dev(self._h, _ptr(scalars_ptr), _ptr(bases_ptr), u64(n), _ptr(out)), "g16_msm_g1_dev")
        return out

    def msm_g2_dev(self, scalars_ptr, bases_ptr, n):
        out = np.zeros(self.g2w, dtype=np.uint64)
        self._check(self._lib.g16_msm_g2_dev(self._h, _ptr(scalars_ptr), _ptr(bases_ptr), u64(n), _ptr(out)), "g16_msm_g2_dev")
        return out

    def ntt_dev(self, data_ptr, n, inverse=False, coset=False, decimation=DIF):
        self._check(self._lib.g16_ntt_dev(self._h, _ptr(data_ptr), u64(n), int(inverse), int(coset), int(decimation)), "g16_ntt_dev")

    def compute_h_dev(self, a_ptr, b_ptr, c_ptr, domain_n):
        self._check(self._lib.g16_compute_h_dev(self._h, _ptr(a_ptr), _ptr(b_ptr), _ptr(c_ptr), u64(domain_n)), "g16_compute_h_dev")

    def compute_h_noalloc_dev(self, a_ptr, b_ptr, c_ptr, n_constraints, domain_n, h_out_ptr):
        """computeH with the library's static device scratch only (no device allocation); domain_n <= 2^16."""
        self._check(self._lib.g16_compute_h_noalloc_dev(self._h, _ptr(a_ptr), _ptr(b_ptr), _ptr(c_ptr), u64(n_constraints),
                                                        u64(domain_n), _ptr(h_out_ptr)), "g16_compute_h_noalloc_dev")

    def g1_fixed_base_mul_dev(self, scalars_ptr, n, out_ptr):
        self._check(self._lib.g16_g1_fixed_base_mul_dev(self._h, _ptr(scalars_ptr), u64(n), _ptr(out_ptr)), "g16_g1_fixed_base_mul_dev")

    def g2_fixed_base_mul_dev(self, scalars_ptr, n, out_ptr):
        self._check(self._lib.g16_g2_fixed_base_mul_dev(self._h, _ptr(scalars_ptr), u64(n), _ptr(out_ptr)), "g16_g2_fixed_base_mul_dev")

    # ---- proving key + prove
    def pk_upload(self, domain_n, g1_a, g1_b, g1_k, g1_z, g2_b, infinity_a, infinity_b, n_wires, k_index=None,
                  shard_index=0, shard_count=1):
        g1_a, g1_b, g1_k, g1_z, g2_b = (_arr(g1_a, self.g1w), _arr(g1_b, self.g1w), _arr(g1_k, self.g1w),
                                        _arr(g1_z, self.g1w), _arr(g2_b, self.g2w))
        ia = np.ascontiguousarray(infinity_a, dtype=np.uint8)
        ib = np.ascontiguousarray(infinity_b, dtype=np.uint8)
        ki = None if k_index is None else np.ascontiguousarray(k_index, dtype=np.uint32)
        d = PkDesc(domain_n, _ptr(g1_a), len(g1_a), _ptr(g1_b), len(g1_b), _ptr(g1_k), len(g1_k), _ptr(g1_z), len(g1_z),
                   _ptr(g2_b), len(g2_b), _ptr(ia), _ptr(ib), n_wires, _ptr(ki), shard_index, shard_count)
        h = u64(0)
        self._check(self._lib.g16_pk_upload(self._h, ctypes.byref(d), ctypes.byref(h)), "g16_pk_upload")
        return h.value

    def pk_upload_from_file(self, fd, domain_n, spans, infinity_a, infinity_b, n_wires, k_index=None,
                            shard_index=0, shard_count=1):
        """spans: {"g1_a": (byte offset, count), "g1_b", "g1_k", "g1_z", "g2_b"} inside the open file `fd`
        (a ProvingKey.WriteDump image); the points never pass through host arrays of the caller."""
        ia = np.ascontiguousarray(infinity_a, dtype=np.uint8)
        ib = np.ascontiguousarray(infinity_b, dtype=np.uint8)
        ki = None if k_index is None else np.ascontiguousarray(k_index, dtype=np.uint32)
        null = ctypes.c_void_p(0)
        key = PkDesc(domain_n, null, spans["g1_a"][1], null, spans["g1_b"][1], null, spans["g1_k"][1], null, spans["g1_z"][1],
                     null, spans["g2_b"][1], _ptr(ia), _ptr(ib), n_wires, _ptr(ki), shard_index, shard_count)
        d = PkFileDesc(key, fd, spans["g1_a"][0], spans["g1_b"][0], spans["g1_k"][0], spans["g1_z"][0], spans["g2_b"][0])
        h = u64(0)
        self._check(self._lib.g16_pk_upload_from_file(self._h, ctypes.byref(d), ctypes.byref(h)), "g16_pk_upload_from_file")
        return h.value

    def pk_wire_range(self, handle):
        """[lo, hi): the wires of W this context's shard of the key reads."""
        lo, hi = u64(0), u64(0)
        self._check(self._lib.g16_pk_wire_range(self._h, u64(handle), ctypes.byref(lo), ctypes.byref(hi)), "g16_pk_wire_range")
        return lo.value, hi.value

    def pk_placement(self, handle, rank=0):
        """[(lo, hi)] x 5 in OUT_* order: the element ranges of G1.A, G1.B, G1.K, G1.Z, G2.B resident on `rank`."""
        lo, hi = (u64 * 5)(), (u64 * 5)()
        self._check(self._lib.g16_pk_placement(self._h, u64(handle), int(rank), lo, hi), "g16_pk_placement")
        return [(int(lo[i]), int(hi[i])) for i in range(5)]

    def pk_free(self, handle):
        self._check(self._lib.g16_pk_free(self._h, u64(handle)), "g16_pk_free")

    def prove(self, pk_handle, W, a, b, c):
        """host arrays in, (5 raw MSM points as a 48-limb array, stats dict) out."""
        W, a, b, c = _arr(W, 4), _arr(a, 4), _arr(b, 4), _arr(c, 4)
        out = np.zeros(self.prove_out_words, dtype=np.uint64)
        st = Stats()
        self._check(self._lib.g16_prove_bn254(self._h, u64(pk_handle), _ptr(W), u64(len(W)), _ptr(a), _ptr(b), _ptr(c),
                                              u64(len(a)), _ptr(out), ctypes.byref(st)), "g16_prove_bn254")
        return out, st.as_dict()

    def prove_dev(self, pk_handle, W_ptr, a_ptr, b_ptr, c_ptr):
        out = np.zeros(self.prove_out_words, dtype=np.uint64)
        st = Stats()
        self._check(self._lib.g16_prove_bn254_dev(self._h, u64(pk_handle), _ptr(W_ptr), _ptr(a_ptr), _ptr(b_ptr), _ptr(c_ptr),
                                                  _ptr(out), ctypes.byref(st)), "g16_prove_bn254_dev")
        return out, st.as_dict()

    # ---- resident bases (precomputed multiples when they fit)
    def bases_upload(self, bases, g2=False):
        b = _arr(bases, self.g2w if g2 else self.g1w)
        h = u64(0)
        self._check(self._lib.g16_bases_upload(self._h, _ptr(b), u64(len(b)), int(g2), 0, ctypes.byref(h)), "g16_bases_upload")
        return h.value

    def bases_upload_dev(self, bases_ptr, n, g2=False):
        h = u64(0)
        self._check(self._lib.g16_bases_upload(self._h, _ptr(bases_ptr), u64(n), int(g2), 1, ctypes.byref(h)), "g16_bases_upload")
        return h.value

    def bases_free(self, handle):
        self._check(self._lib.g16_bases_free(self._h, u64(handle)), "g16_bases_free")

    def msm_resident_dev(self, handle, scalars_ptr, g2=False):
        out = np.zeros(self.g2w if g2 else self.g1w, dtype=np.uint64)
        self._check(self._lib.g16_msm_resident_dev(self._h, u64(handle), _ptr(scalars_ptr), _ptr(out)), "g16_msm_resident_dev")
        return out

    def msm_resident(self, handle, scalars, g2=False):
        s = _arr(scalars, 4) if len(scalars) else np.zeros((0, 4), dtype=np.uint64)
        out = np.zeros(self.g2w if g2 else self.g1w, dtype=np.uint64)
        self._check(self._lib.g16_msm_resident(self._h, u64(handle), _ptr(s), _ptr(out)), "g16_msm_resident")
        return out

    # ---- BSB22 commitment keys
    def commitment_key_upload(self, basis, basis_exp_sigma):
        empty = np.zeros((0, self.g1w), dtype=np.uint64)
        b = _arr(basis, self.g1w) if len(basis) else empty
        e = _arr(basis_exp_sigma, self.g1w) if len(basis_exp_sigma) else empty
        if len(b) != len(e):
            raise G16Error("commitment key: Basis and BasisExpSigma differ in length")
        h = u64(0)
        self._check(self._lib.g16_commitment_key_upload(self._h, _ptr(b), _ptr(e), u64(len(b)), ctypes.byref(h)),
                    "g16_commitment_key_upload")
        return h.value

    def commitment_key_free(self, handle):
        self._check(self._lib.g16_commitment_key_free(self._h, u64(handle)), "g16_commitment_key_free")

    def commit(self, handle, values, want_pok=True):
        """(commitment, pok) = (sum v_j Basis_j, sum v_j BasisExpSigma_j); pok is None unless want_pok."""
        v = _arr(values, 4) if len(values) else np.zeros((0, 4), dtype=np.uint64)
        cm = np.zeros(self.g1w, dtype=np.uint64)
        pok = np.zeros(self.g1w, dtype=np.uint64) if want_pok else None
        self._check(self._lib.g16_commit(self._h, u64(handle), _ptr(v), u64(len(v)), _ptr(cm), _ptr(pok)), "g16_commit")
        return cm, pok

    # ---- constraint matrices on the device
    def r1cs_upload(self, n_constraints, n_wires, coeffs, matrices):
        """matrices: three (row_ptr, wire, coeff_id) uint32 triples for L, R, O (CSR over the constraints)."""
        coeffs = _arr(coeffs, 4) if len(coeffs) else np.zeros((0, 4), dtype=np.uint64)
        keep = []
        d = R1csDesc()
        d.n_constraints, d.n_wires, d.coeffs, d.n_coeffs = n_constraints, n_wires, coeffs.ctypes.data, len(coeffs)
        for m, (rp, wi, ci) in enumerate(matrices):
            rp, wi, ci = (np.ascontiguousarray(x, dtype=np.uint32) for x in (rp, wi, ci))
            assert len(rp) == n_constraints + 1 and len(wi) == len(ci) == int(rp[-1]), (len(rp), len(wi), len(ci))
            keep += [rp, wi, ci]
            d.row_ptr[m], d.wire[m], d.coeff_id[m] = rp.ctypes.data, wi.ctypes.data, ci.ctypes.data
        h = u64(0)
        self._check(self._lib.g16_r1cs_upload(self._h, ctypes.byref(d), ctypes.byref(h)), "g16_r1cs_upload")
        return h.value

    def r1cs_free(self, handle):
        self._check(self._lib.g16_r1cs_free(self._h, u64(handle)), "g16_r1cs_free")

    def r1cs_eval_dev(self, handle, W_ptr, a_ptr, b_ptr, c_ptr, domain_n):
        self._check(self._lib.g16_r1cs_eval_dev(self._h, u64(handle), _ptr(W_ptr), _ptr(a_ptr), _ptr(b_ptr), _ptr(c_ptr),
                                                u64(domain_n)), "g16_r1cs_eval_dev")

    def prove_wires(self, pk_handle, r1cs_handle, W):
        """only the wire vector crosses the boundary: a, b, c are evaluated on the device from the resident matrices."""
        W = _arr(W, 4)
        out = np.zeros(self.prove_out_words, dtype=np.uint64)
        st = Stats()
        self._check(self._lib.g16_prove_bn254_wires(self._h, u64(pk_handle), u64(r1cs_handle), _ptr(W), u64(len(W)), _ptr(out),
                                                    ctypes.byref(st)), "g16_prove_bn254_wires")
        return out, st.as_dict()

    # ---- distributed computeH building blocks
    def dist_ntt_stages_dev(self, data_ptr, count, n_global, world, rank, table, dif, cyclic, split):
        self._check(self._lib.g16_dist_ntt_stages_dev(self._h, _ptr(data_ptr), int(count), u64(n_global), int(world), int(rank),
                                                      int(table), int(dif), int(cyclic), int(split)), "g16_dist_ntt_stages_dev")

    def dist_permute_dev(self, src_ptr, dst_ptr, count, n_local, world, block_layout, to_wire):
        self._check(self._lib.g16_dist_permute_dev(self._h, _ptr(src_ptr), _ptr(dst_ptr), int(count), u64(n_local), int(world),
                                                   int(block_layout), int(to_wire)), "g16_dist_permute_dev")

    def dist_pointwise_h_dev(self, abc_ptr, n_local, n_global):
        self._check(self._lib.g16_dist_pointwise_h_dev(self._h, _ptr(abc_ptr), u64(n_local), u64(n_global)), "g16_dist_pointwise_h_dev")

    def prove_msms_begin_dev(self, pk_handle, W_ptr):
        """start the four MSMs over W; they run while the caller computes h; finish with prove_msms_dev."""
        self._check(self._lib.g16_prove_msms_begin_dev(self._h, u64(pk_handle), _ptr(W_ptr)), "g16_prove_msms_begin_dev")

    def prove_msms_dev(self, pk_handle, W_ptr, h_local_ptr):
        out = np.zeros(self.prove_out_words, dtype=np.uint64)
        st = Stats()
        self._check(self._lib.g16_prove_msms_dev(self._h, u64(pk_handle), _ptr(W_ptr), _ptr(h_local_ptr), _ptr(out), ctypes.byref(st)),
                    "g16_prove_msms_dev")
        return out, st.as_dict()

    def imad_peak(self):
        a, b = ctypes.c_double(0), ctypes.c_double(0)
        self._check(self._lib.g16_imad_peak(self._h, ctypes.byref(a), ctypes.byref(b)), "g16_imad_peak")
        return a.value, b.value


def _pipe_rates(self):
    out = (ctypes.c_double * 3)()
    self._check(self._lib.g16_pipe_rates(self._h, out), "g16_pipe_rates")
    return out[0], out[1], out[2]


Context.pipe_rates = _pipe_rates


def device_count(curve="bn254"):
    return int(load(curve).g16_device_count())


def kernel_launch_count(curve="bn254"):
    return int(load(curve).g16_kernel_launch_count())


# ---- host-only helpers (no GPU needed)
def g1_add(a, b, curve="bn254"):
    w = CURVES[curve][1]
    a, b = _arr(a, w), _arr(b, w)
    out = np.zeros(w, dtype=np.uint64)
    assert load(curve).g16_g1_add(_ptr(a), _ptr(b), _ptr(out)) == 0
    return out


def g2_add(a, b, curve="bn254"):
    w = CURVES[curve][2]
    a, b = _arr(a, w), _arr(b, w)
    out = np.zeros(w, dtype=np.uint64)
    assert load(curve).g16_g2_add(_ptr(a), _ptr(b), _ptr(out)) == 0
    return out


def g1_scalar_mul(a, k, curve="bn254"):
    w = CURVES[curve][1]
    a, k = _arr(a, w), _arr(k, 4)
    out = np.zeros(w, dtype=np.uint64)
    assert load(curve).g16_g1_scalar_mul(_ptr(a), _ptr(k), _ptr(out)) == 0
    return out


def g2_scalar_mul(a, k, curve="bn254"):
    w = CURVES[curve][2]
    a, k = _arr(a, w), _arr(k, 4)
    out = np.zeros(w, dtype=np.uint64)
    assert load(curve).g16_g2_scalar_mul(_ptr(a), _ptr(k), _ptr(out)) == 0
    return out


def assemble_proof(pk_points, r, s, msm_points, curve="bn254"):
    """pk_points: alpha, beta, delta (G1), beta, delta (G2) = 56 limbs for BN254; returns Ar||Bs||Krs (32 limbs)."""
    _, w1, w2 = CURVES[curve]
    pk_points = np.ascontiguousarray(pk_points, dtype=np.uint64).reshape(-1)
    msm_points = np.ascontiguousarray(msm_points, dtype=np.uint64).reshape(-1)
    assert pk_points.size == 3 * w1 + 2 * w2 and msm_points.size == 4 * w1 + w2, (pk_points.size, msm_points.size)
    r, s = _arr(r, 4), _arr(s, 4)
    out = np.zeros(2 * w1 + w2, dtype=np.uint64)
    assert load(curve).g16_assemble_proof_bn254(_ptr(pk_points), _ptr(r), _ptr(s), _ptr(msm_points), _ptr(out)) == 0
    return out
