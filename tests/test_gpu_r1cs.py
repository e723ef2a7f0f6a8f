"""GPU: a = L.w, b = R.w, c = O.w evaluated on the device from resident constraint matrices --
bit-exact against the oracle's solver outputs, and proofs through the wires-only entry point identical to the ones
through g16_prove_bn254 (which receives a, b, c from the host)."""
import numpy as np
import pytest

from gnark_icicle_b200 import lib
import gnark_icicle_b200 as g16
from oracle import bn254 as C
from oracle import cpu as ocpu
from oracle import groth16 as og
from oracle import r1cs as orcs
from oracle.prng import Xoshiro, SEED_BASE

from helpers import ChainCase, chain_csr, csr_from_oracle, pk_from_python, assert_limbs_equal

pytestmark = pytest.mark.gpu
R = C.R


def eval_on_device(ctx, cs, W, domain_n):
    import torch
    h = cs.setup_device(ctx)
    Wd = torch.from_numpy(np.ascontiguousarray(W).view(np.int64)).cuda()
    abc = torch.full((3, domain_n, 4), -1, dtype=torch.int64, device="cuda")   # poison: padding must be written
    torch.cuda.synchronize()
    ctx.r1cs_eval_dev(h, Wd.data_ptr(), abc[0].data_ptr(), abc[1].data_ptr(), abc[2].data_ptr(), domain_n)
    return abc.cpu().numpy().view(np.uint64)


@pytest.mark.parametrize("case", ["cubic", "chain"])
def test_eval_matches_python_oracle(ctx, case):
    rng = Xoshiro(SEED_BASE + 80)
    if case == "cubic":
        cs_py, w = orcs.cubic_circuit(), orcs.cubic_witness(3, 35)
    else:
        cs_py, _ = orcs.mimc_chain(301, rng)
        w = orcs.mimc_chain_witness(cs_py, rng.field(R), rng.field(R))
        w[1] = sum(cf * w[i] for cf, i in cs_py.constraints[-1][0]) % R
    a, b, c = cs_py.abc(w)
    cs = csr_from_oracle(cs_py, ocpu.fr_arr)
    n = 4 if case == "cubic" else 512
    try:
        got = eval_on_device(ctx, cs, ocpu.fr_arr(w), n)
        for m, want in enumerate((a, b, c)):
            assert_limbs_equal(got[m, :len(want)], ocpu.fr_arr(want), f"matrix {m}")
            assert not got[m, len(want):].any(), "padding rows must be zero"
    finally:
        cs.free_device()


@pytest.mark.parametrize("nc", [4, 1000, 1 << 20])
def test_chain_eval_matches_c_oracle_and_proves_identically(ctx, nc):
    makers = {} if nc <= 1000 else dict(g1_mul=ctx.g1_fixed_base_mul, g2_mul=ctx.g2_fixed_base_mul)
    case = ChainCase(nc, seed=31, **makers)
    cs = chain_csr(nc, SEED_BASE + 2000 + 31)
    pk = case.pk
    try:
        pk.setup_device_pointers()
        ctx = pk.device[0]
        got = eval_on_device(ctx, cs, case.W, case.n)
        for m, want in enumerate((case.a, case.b, case.c)):
            assert_limbs_equal(got[m, :nc], want, f"matrix {m}")
            assert not got[m, nc:].any()
        sol = g16.R1CSSolution(W=case.W, A=case.a, B=case.b, C=case.c)
        want_pts, _ = g16.prove_msms(pk, sol)
        got_pts, st = g16.prove_msms_wires(pk, cs, case.W)
        assert np.array_equal(got_pts, want_pts)
        assert st["h2d_bytes"] == case.nw * 32
        proof = g16.Prove(cs, pk, case.W, g16.WithIcicleAcceleration(), g16.WithRandomness(case.r, case.s))
        assert proof.raw() == g16.assemble(pk, want_pts, case.r, case.s).raw()
    finally:
        cs.free_device()
        pk.free_device()


def test_upload_rejects_malformed_matrices(ctx):
    one = ocpu.fr_arr([1])
    ok = (np.array([0, 1], dtype=np.uint32), np.array([0], dtype=np.uint32), np.array([0], dtype=np.uint32))
    bad_wire = (ok[0], np.array([7], dtype=np.uint32), ok[2])
    bad_coeff = (ok[0], ok[1], np.array([3], dtype=np.uint32))
    for mats in ((bad_wire, ok, ok), (ok, bad_coeff, ok)):
        with pytest.raises(lib.G16Error):
            ctx.r1cs_upload(1, 2, one, mats)
    h = ctx.r1cs_upload(1, 2, one, (ok, ok, ok))
    ctx.r1cs_free(h)
    with pytest.raises(lib.G16Error):
        ctx.r1cs_free(h)
