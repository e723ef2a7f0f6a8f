"""GPU parity of the multi-GPU path, emulated on one GPU: the four-step (cyclic/block + all-to-all) computeH over
2, 4 and 8 virtual ranks must reproduce the oracle's h bit for bit, and the sharded MSM-only prove fed with each
rank's block of h must combine to the single-GPU result."""
import numpy as np
import pytest
import torch

from helpers import ChainCase, fr_rand, assert_limbs_equal
from oracle import cpu as ocpu
import gnark_icicle_b200 as g16
from gnark_icicle_b200 import dist as gdist, lib

pytestmark = pytest.mark.gpu


def _parts(a, b, c, n, world):
    out = []
    for r in range(world):
        abc = np.stack([gdist.cyclic_slice(x, n, r, world) for x in (a, b, c)])
        out.append(torch.from_numpy(abc.view(np.int64)).cuda())
    return out


@pytest.mark.parametrize("logn,world", [(4, 2), (6, 8), (12, 2), (12, 4), (16, 8), (18, 4)])
def test_distributed_compute_h_matches_oracle(ctx, logn, world):
    n = 1 << logn
    nc = n - 3
    a, b = fr_rand(300 + logn, nc), fr_rand(301 + logn, nc)
    c = ocpu.fr_mul_vec(a, b)
    c[1::3] = fr_rand(302 + logn, nc)[1::3]
    blocks = gdist.compute_h_emulated(ctx, _parts(a, b, c, n, world), n)
    got = np.concatenate([blk.cpu().numpy().view(np.uint64) for blk in blocks])
    assert_limbs_equal(got, ocpu.compute_h(a, b, c, n), f"distributed computeH n=2^{logn} world={world}")


def test_world_one_degenerates_to_local(ctx):
    n = 1 << 10
    a, b, c = fr_rand(310, n), fr_rand(311, n), fr_rand(312, n)
    blocks = gdist.compute_h_emulated(ctx, _parts(a, b, c, n, 1), n)
    assert_limbs_equal(blocks[0].cpu().numpy().view(np.uint64), ocpu.compute_h(a, b, c, n), "world=1")


def test_sharded_prove_with_distributed_h(ctx):
    world = 4
    case = ChainCase(5000, seed=9)
    sol = g16.R1CSSolution(W=case.W, A=case.a, B=case.b, C=case.c)
    whole, _ = g16.prove_msms(case.pk, sol)
    case.pk.free_device()
    blocks = gdist.compute_h_emulated(ctx, _parts(case.a, case.b, case.c, case.n, world), case.n)
    W_d = torch.from_numpy(case.W.view(np.int64)).cuda()
    pk = case.pk
    outs = []
    for r in range(world):
        h = ctx.pk_upload(pk.domain_n, pk.g1_a, pk.g1_b, pk.g1_k, pk.g1_z, pk.g2_b, pk.infinity_a, pk.infinity_b, pk.nb_wires,
                          shard_index=r, shard_count=world)
        pts, st = ctx.prove_msms_dev(h, W_d.data_ptr(), blocks[r].data_ptr())
        # the two-call form (the W MSMs started first, h handed in later) gives the same points
        ctx.prove_msms_begin_dev(h, W_d.data_ptr())
        with pytest.raises(lib.G16Error, match="has not been finished"):
            ctx.prove_msms_begin_dev(h, W_d.data_ptr())
        pts2, _ = ctx.prove_msms_dev(h, W_d.data_ptr(), blocks[r].data_ptr())
        assert np.array_equal(pts, pts2)
        outs.append(pts)
        ctx.pk_free(h)
    assert_limbs_equal(g16.combine_shards(outs), whole, "sharded MSMs over distributed h")
