import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, time
from helpers import ChainCase
import gnark_icicle_b200 as g16
from gnark_icicle_b200 import lib
ctx0 = lib.Context(0)
for nc in (1<<15, 1<<18):
    case = ChainCase(nc, seed=3, g1_mul=ctx0.g1_fixed_base_mul, g2_mul=ctx0.g2_fixed_base_mul)
    pk = case.pk
    t=time.time(); pk.setup_device_pointers(); print("upload+precompute s", time.time()-t)
    ctx = pk.device[0]
    sol = g16.R1CSSolution(W=case.W, A=case.a, B=case.b, C=case.c)
    p1, st1 = g16.prove_msms(pk, sol)
    ctx.set("msm_precompute", 0)
    p0, st0 = g16.prove_msms(pk, sol)
    ctx.set("msm_precompute", 1)
    print(nc, "equal:", np.array_equal(p0,p1), st1["msm_c"], st1["msm_windows"], st0["msm_c"], st0["msm_windows"], st1["ms_total"], st0["ms_total"])
    pk.free_device()
