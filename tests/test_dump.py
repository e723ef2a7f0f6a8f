"""ProvingKey.WriteDump / ReadDump and the direct device load.  The CPU tests check the dump
reader and writer of the product against the oracle's independent writer (marshal.go:386-550 layout; the reference
holds no dump fixture, icicle/marshal_test.go:27-67 only checks that icicle and native keys serialise alike); the GPU
tests prove with a key that went file -> HBM and compare with the key uploaded from host arrays."""
import io
import os

import numpy as np
import pytest

import gnark_icicle_b200 as g16
from gnark_icicle_b200 import dump, lib
from oracle import groth16 as G, r1cs as RC, curve as C, dump as odump
from oracle import cpu as ocpu
from oracle.prng import Xoshiro

from helpers import pk_from_python


def _py_case(ncm, seed=3, nc=40):
    rng = Xoshiro(seed)
    cs = RC.committed_chain(nc, rng, ncm) if ncm else RC.mimc_chain(nc, rng)[0]
    pk_py, vk, scal = G.setup(cs, rng)
    pk = pk_from_python(pk_py, cs.nb_wires)
    g1 = lambda pts: ocpu.g1_arr(pts) if len(pts) else np.zeros((0, 8), dtype=np.uint64)
    pk.commitment_keys = [g16.bsb22.CommitmentKey(g1(k.basis), g1(k.basis_exp_sigma)) for k in pk_py.commitment_keys]
    return rng, cs, pk_py, pk


@pytest.mark.parametrize("ncm", [0, 2, 3])
def test_write_dump_matches_oracle_writer_and_round_trips(tmp_path, ncm):
    rng, cs, pk_py, pk = _py_case(ncm)
    want = odump.write_dump(pk_py)
    assert dump.dump_bytes(pk) == want
    path = tmp_path / "pk.dump"
    path.write_bytes(want)
    back = dump.read_dump(str(path), nb_public=cs.nb_public)
    for name in ("g1_alpha", "g1_beta", "g1_delta", "g2_beta", "g2_delta", "g1_a", "g1_b", "g1_z", "g1_k", "g2_b",
                 "infinity_a", "infinity_b"):
        assert np.array_equal(getattr(back, name), getattr(pk, name)), name
    assert back.domain_n == pk.domain_n and back.nb_wires == pk.nb_wires
    assert len(back.commitment_keys) == ncm
    for a, b in zip(back.commitment_keys, pk.commitment_keys):
        assert np.array_equal(a.basis, b.basis) and np.array_equal(a.basis_exp_sigma, b.basis_exp_sigma)
    with open(path, "rb") as f:
        lay = dump.read_dump_layout(f)
    off, cnt = lay.spans["g1_z"]
    assert cnt == pk.domain_n - 1 and want[off:off + 64] == pk.g1_z[0].tobytes()


def test_dump_error_paths(tmp_path):
    rng, cs, pk_py, pk = _py_case(0)
    good = dump.dump_bytes(pk)
    with pytest.raises(lib.G16Error, match="read marker"):
        dump.read_dump_layout(io.BytesIO(b"\x00" * 8 + good[8:]))
    with pytest.raises((lib.G16Error, Exception)):
        dump.read_dump_layout(io.BytesIO(good[:len(good) - 100]))


@pytest.mark.gpu
@pytest.mark.parametrize("shards", [1, 3])
def test_device_load_from_dump_proves_like_host_upload(tmp_path, shards):
    from helpers import ChainCase
    case = ChainCase(20000, seed=4)          # above the precompute threshold: tables are built from the loaded rows
    path = tmp_path / "pk.dump"
    with open(path, "wb") as f:
        dump.write_dump(case.pk, f)
    sol = g16.R1CSSolution(W=case.W, A=case.a, B=case.b, C=case.c)
    want, _ = g16.prove_msms(case.pk, sol)
    case.pk.free_device()
    blocks = []
    for si in range(shards):
        pk = dump.read_dump_to_device(str(path), nb_public=2, cfg=g16.NewProverConfig(g16.WithDevice(0, si, shards)))
        assert pk.g1_a.shape == (0, 8) and pk.device is not None
        pts, _ = g16.prove_msms(pk, sol)
        blocks.append(pts)
        pk.free_device()
    got = g16.combine_shards(blocks)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_device_load_error_paths(tmp_path):
    rng, cs, pk_py, pk = _py_case(0)
    path = tmp_path / "pk.dump"
    path.write_bytes(dump.dump_bytes(pk)[:-64])      # last slice is cut short
    with pytest.raises(lib.G16Error):
        dump.read_dump_to_device(str(path))
    ctx = lib.Context(0)
    with open(path, "rb") as f:
        lay = dump.read_dump_layout(io.BytesIO(dump.dump_bytes(pk)))
        spans = dict(lay.spans)
        spans["g2_b"] = (spans["g2_b"][0] + 64, spans["g2_b"][1])       # runs past the end of the file
        with pytest.raises(lib.G16Error, match="shorter than its slice headers"):
            ctx.pk_upload_from_file(f.fileno(), lay.domain_n, spans, lay.infinity_a, lay.infinity_b, lay.nb_wires)
    ctx.close()
