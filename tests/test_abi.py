"""CPU-only: the C-ABI library loads, exports every symbol include/g16b200.h declares, its host-side helpers
(epilogue arithmetic) agree with the oracle, and the accelerated path refuses to run without a GPU."""
import json
import os
import re

import numpy as np
import pytest

from gnark_icicle_b200 import lib
import gnark_icicle_b200 as g16
from oracle import bn254 as C
from oracle import cpu as ocpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def limbs(hexstr, width):
    return np.frombuffer(bytes.fromhex(hexstr), dtype=np.uint64).reshape(-1, width).copy()


def test_library_exports_every_declared_symbol():
    with open(os.path.join(ROOT, "include", "g16b200.h")) as f:
        hdr = f.read()
    declared = set(re.findall(r"\b(g16_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations found"
    handle = lib.load()
    missing = [s for s in sorted(declared) if not hasattr(handle, s)]
    assert not missing, missing
    assert declared == set(lib.SYMBOLS), declared ^ set(lib.SYMBOLS)


def test_header_cites_reference_interfaces():
    with open(os.path.join(ROOT, "include", "g16b200.h")) as f:
        hdr = f.read()
    for cite in ("icicle.go:31-130", "icicle.go:302", "icicle.go:382", "icicle.go:489", "prove.go:356-399", "setup.go:251"):
        assert cite in hdr, cite


def test_stats_struct_layout_matches_header():
    # 4+4+4+20+4 = 36 | 20+20 = 76 -> pad 80 | 40 = 120 | 16 = 136 | 8 = 144 | 24 = 168
    import ctypes
    assert ctypes.sizeof(lib.Stats) == 168
    assert lib.Stats.msm_n.offset == 80 and lib.Stats.ms_accumulate.offset == 144


def test_host_group_helpers_match_oracle():
    g1 = ocpu.g1_arr([C.G1_GEN])[0]
    g2 = ocpu.g2_arr([C.G2_GEN])[0]
    k = ocpu.fr_random(3, 4)
    p = lib.g1_scalar_mul(g1, k[0])
    assert np.array_equal(p, ocpu.g1_mul(g1, k[0]))
    q = lib.g2_scalar_mul(g2, k[1])
    assert np.array_equal(q, ocpu.g2_mul(g2, k[1]))
    assert np.array_equal(lib.g1_add(p, g1), ocpu.g1_add(p, g1))
    assert np.array_equal(lib.g2_add(q, g2), ocpu.g2_add(q, g2))
    # identities: P + inf, P + P (doubling path), P + (-P)
    zero1 = np.zeros(8, dtype=np.uint64)
    assert np.array_equal(lib.g1_add(p, zero1), p) and np.array_equal(lib.g1_add(zero1, p), p)
    assert np.array_equal(lib.g1_add(p, p), ocpu.g1_add(p, p))
    neg = ocpu.g1_arr([C.g1_neg(ocpu.g1_pt(p))])[0]
    assert not lib.g1_add(p, neg).any()
    assert not lib.g1_scalar_mul(p, np.zeros(4, dtype=np.uint64)).any()


def test_assemble_matches_golden_cubic_proof():
    """host epilogue (prove.go:180-302) on the golden MSM points must reproduce the golden proof bytes."""
    with open(os.path.join(GOLD, "cubic_proof.json")) as f:
        g = json.load(f)
    pkp = np.concatenate([limbs(g[k], 8).reshape(-1) for k in ("g1_alpha", "g1_beta", "g1_delta")]
                         + [limbs(g[k], 16).reshape(-1) for k in ("g2_beta", "g2_delta")])
    pts = np.frombuffer(bytes.fromhex(g["msm_points"]), dtype=np.uint64)
    out = lib.assemble_proof(pkp, limbs(g["r"], 4)[0], limbs(g["s"], 4)[0], pts)
    assert out.tobytes().hex() == g["proof"]


def test_combine_shards_is_group_addition():
    k = ocpu.fr_random(9, 10)
    g1 = ocpu.g1_mul_gen_batch(k[:8])
    g2 = ocpu.g2_mul_gen_batch(k[8:10])
    blk0 = np.concatenate([g1[0], g1[1], g1[2], g1[3], g2[0]])
    blk1 = np.concatenate([g1[4], g1[5], g1[6], g1[7], g2[1]])
    got = g16.combine_shards([blk0, blk1])
    want = np.concatenate([ocpu.g1_add(g1[i], g1[i + 4]) for i in range(4)] + [ocpu.g2_add(g2[0], g2[1])])
    assert np.array_equal(got, want)


def test_no_cpu_fallback():
    """without a CUDA device the context cannot be created and Prove raises -- nothing routes to a CPU path."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: the failure mode under test needs a GPU-less machine")
    with pytest.raises(lib.G16Error):
        lib.Context(0)
    with pytest.raises(lib.G16Error):
        g16.Prove(None, None, None)          # missing WithIcicleAcceleration()


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "gnark_icicle_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert "import oracle" not in src and "from oracle" not in src and "oracle/" not in src.replace("oracle/ ", ""), fn
