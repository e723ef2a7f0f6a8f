"""One resident-bases MSM of a given size, as a rank of a multi-GPU context sees its piece of a key array: for launch
lists (ncu --metrics gpu__time_duration.sum) of the kernels that do not shrink with the piece (sort heads, inversion
trees, bucket reduction).

    python tools/prof_piece.py [--n 900000] [--g2] [--runs 2]
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gnark_icicle_b200 import lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=900000)
ap.add_argument("--g2", action="store_true")
ap.add_argument("--runs", type=int, default=2)
ap.add_argument("--ba", type=int, nargs="*", default=[-1], help="msm_ba_rounds values to try (-1 = library default)")
ap.add_argument("--pre-c", type=int, nargs="*", default=[0], help="window bits of the table (0 = the library's choice); several values: one upload each")
args = ap.parse_args()
dev = torch.device("cuda:0")
ctx = lib.Context(0)
gen = torch.Generator(device=dev)
gen.manual_seed(5)


def rnd(n):
    x = torch.randint(-(2 ** 63), 2 ** 63 - 1, (n, 4), dtype=torch.int64, device=dev, generator=gen)
    x[:, 3] &= 0x0FFFFFFFFFFFFFFF
    return x


k, s = rnd(args.n), rnd(args.n)
w = ctx.g2w if args.g2 else ctx.g1w
bases = torch.empty((args.n, w), dtype=torch.int64, device=dev)
torch.cuda.synchronize()
(ctx.g2_fixed_base_mul_dev if args.g2 else ctx.g1_fixed_base_mul_dev)(k.data_ptr(), args.n, bases.data_ptr())
ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
for c in args.pre_c:
    ctx.set("msm_pre_c", c)
    h = ctx.bases_upload_dev(bases.data_ptr(), args.n, g2=args.g2)
    for ba in args.ba:
        if ba >= 0:
            ctx.set("msm_ba_rounds", ba)
        best = 1e9
        for _ in range(args.runs):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext)
            ctx.msm_resident_dev(h, s.data_ptr(), g2=args.g2)
            e1.record(ext)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        print("MSM_MS", "n", args.n, "g2" if args.g2 else "g1", "pre_c", c, "ba_rounds", ba, best)
    ctx.bases_free(h)
