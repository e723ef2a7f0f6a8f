"""times a 2^19-point G1 MSM for every window size c (device-resident inputs)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnark_icicle_b200 import lib
ctx = lib.Context(0)
ext = torch.cuda.ExternalStream(ctx.stream)
m = 1 << 19
k = torch.randint(0, 2**62, (m, 4), dtype=torch.int64, device="cuda"); k[:, 3] &= 0x0FFFFFFFFFFFFFFF
bases = torch.empty((m, 8), dtype=torch.int64, device="cuda")
ctx.g1_fixed_base_mul_dev(k.data_ptr(), m, bases.data_ptr())
s = torch.randint(0, 2**62, (m, 4), dtype=torch.int64, device="cuda"); s[:, 3] &= 0x0FFFFFFFFFFFFFFF
only = [int(sys.argv[1])] if len(sys.argv) > 1 else range(8, 17)
for c in only:
    ctx.set("msm_c", c)
    ctx.msm_g1_dev(s.data_ptr(), bases.data_ptr(), m)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext); r = ctx.msm_g1_dev(s.data_ptr(), bases.data_ptr(), m); e1.record(ext); torch.cuda.synchronize()
    print(c, round(e0.elapsed_time(e1), 3), "ms", hex(int(r[0])))
