"""resident-bases G1 MSM time against the window of the table: python experiments/pre_c_sweep.py <c> (g16_ctx_set "msm_pre_c")."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gnark_icicle_b200 import lib
ctx = lib.Context(0)
C = int(sys.argv[1]) if len(sys.argv) > 1 else 0
ctx.set("msm_pre_c", C)
gen = torch.Generator(device="cuda"); gen.manual_seed(1)
ext = torch.cuda.ExternalStream(ctx.stream)
for logn in (17, 18, 19, 20, 21):
    m = 1 << logn
    k = torch.randint(-(2**63), 2**63 - 1, (m, 4), dtype=torch.int64, device="cuda", generator=gen); k[:, 3] &= 0x0FFFFFFFFFFFFFFF
    s = torch.randint(-(2**63), 2**63 - 1, (m, 4), dtype=torch.int64, device="cuda", generator=gen); s[:, 3] &= 0x0FFFFFFFFFFFFFFF
    bases = torch.empty((m, 8), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    ctx.g1_fixed_base_mul_dev(k.data_ptr(), m, bases.data_ptr())
    h = ctx.bases_upload_dev(bases.data_ptr(), m)
    for _ in range(3):
        ctx.msm_resident_dev(h, s.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(5):
        ctx.msm_resident_dev(h, s.data_ptr())
    e1.record(ext)
    torch.cuda.synchronize()
    print(f"c={C or 'auto'} 2^{logn}: {e0.elapsed_time(e1) / 5:.3f} ms")
    ctx.bases_free(h)
