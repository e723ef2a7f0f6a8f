"""Profiling target: a few BN254 proves at 2^log_n constraints on a synthetic key made on the device (random bases from the
fixed-base kernel, MiMC-chain infinity shape), MSMs serialised (streams=1) so that per-kernel times are meaningful.

    python tools/prof_prove.py [--log-n 22] [--proves 2] [--streams 1]
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/prof_prove.py
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np   # noqa: E402
import torch         # noqa: E402
import bench         # noqa: E402
from gnark_icicle_b200 import lib   # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--log-n", type=int, default=22)
ap.add_argument("--proves", type=int, default=2)
ap.add_argument("--streams", type=int, default=1)
ap.add_argument("--curve", default="bn254")
ap.add_argument("--ba", type=int, default=-1, help="msm_ba_rounds_g1 (-1: library default)")
ap.add_argument("--ba2", type=int, default=-1, help="msm_ba_rounds_g2 (-1: library default)")
args = ap.parse_args()
dev = torch.device("cuda:0")
ctx = lib.Context(0, curve=args.curve)
nc = 1 << args.log_n
key, sd, sh, n, nw = bench.build_inputs(torch, ctx, nc, dev)
h = ctx.pk_upload(key["domain_n"], key["g1_a"], key["g1_b"], key["g1_k"], key["g1_z"], key["g2_b"], key["inf_a"], key["inf_b"], nw)
del key
ctx.set("streams", args.streams)
if args.ba >= 0:
    ctx.set("msm_ba_rounds_g1", args.ba)
if args.ba2 >= 0:
    ctx.set("msm_ba_rounds_g2", args.ba2)
work = [torch.empty_like(sd[k]) for k in ("a", "b", "c")]
for it in range(args.proves):
    for w, k in zip(work, ("a", "b", "c")):
        w.copy_(sd[k])
    torch.cuda.synchronize()
    pts, st = ctx.prove_dev(h, sd["W"].data_ptr(), work[0].data_ptr(), work[1].data_ptr(), work[2].data_ptr())
print("PROVE_MS", st["ms_total"], "compute_h", st["ms_compute_h"], "msm", [round(x, 3) for x in st["ms_msm"]],
      "acc", [round(x, 3) for x in st["ms_accumulate"]], "launches", st["kernel_launches"])
