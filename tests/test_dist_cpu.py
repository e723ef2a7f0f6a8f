"""CPU-only, world_size 2 over gloo: the N>1 host logic of bench.py / Prove -- every rank owns a point-range
shard, the 384-byte result blocks are all-gathered and rank 0 adds them (level 2)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import gnark_icicle_b200 as g16
    from oracle import cpu as ocpu
    # the shard's partial MSM results, here produced by the CPU oracle over this rank's point range
    n = 64
    k = ocpu.fr_random(21, n)
    s = ocpu.fr_random(22, n)
    pts, pts2 = ocpu.g1_mul_gen_batch(k), ocpu.g2_mul_gen_batch(k)
    lo, hi = n * rank // world, n * (rank + 1) // world
    part1 = ocpu.msm_g1(pts[lo:hi], s[lo:hi])
    part2 = ocpu.msm_g2(pts2[lo:hi], s[lo:hi])
    blk = np.concatenate([part1, part1, part1, part1, part2])
    t = torch.from_numpy(blk.view(np.int64).copy())
    gathered = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(gathered, t)
    if rank == 0:
        total = g16.combine_shards([g.numpy().view(np.uint64) for g in gathered])
        ok = np.array_equal(total[:8], ocpu.msm_g1(pts, s)) and np.array_equal(total[32:48], ocpu.msm_g2(pts2, s))
        q.put(bool(ok))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_results_combine_over_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
