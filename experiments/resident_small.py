"""per-kernel times of the resident (precomputed) MSM at small sizes -- run under ncu for a launch list."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gnark_icicle_b200 import lib
ctx = lib.Context(0)
gen = torch.Generator(device="cuda"); gen.manual_seed(1)
for logn in (16, 18):
    m = 1 << logn
    k = torch.randint(-(2**63), 2**63 - 1, (m, 4), dtype=torch.int64, device="cuda", generator=gen); k[:, 3] &= 0x0FFFFFFFFFFFFFFF
    s = torch.randint(-(2**63), 2**63 - 1, (m, 4), dtype=torch.int64, device="cuda", generator=gen); s[:, 3] &= 0x0FFFFFFFFFFFFFFF
    bases = torch.empty((m, 8), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    ctx.g1_fixed_base_mul_dev(k.data_ptr(), m, bases.data_ptr())
    h = ctx.bases_upload_dev(bases.data_ptr(), m)
    a = ctx.msm_resident_dev(h, s.data_ptr())
    b = ctx.msm_g1_dev(s.data_ptr(), bases.data_ptr(), m)
    print(logn, np.array_equal(a, b))
