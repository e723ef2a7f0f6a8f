"""tunes the NTT tile shape: prints ms for a 2^22 / 2^24 coset DIT for the tile shape compiled into ntt.cu (ntt_log_tile / ntt_log_cols: edit, rebuild, rerun)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnark_icicle_b200 import lib
ctx = lib.Context(0)
ext = torch.cuda.ExternalStream(ctx.stream)
out = {}
for logn in (22, 24):
    m = 1 << logn
    x = torch.randint(0, 2**62, (m, 4), dtype=torch.int64, device="cuda")
    x[:, 3] &= 0x0FFFFFFFFFFFFFFF
    ctx.ntt_dev(x.data_ptr(), m, False, True, 1)
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext); ctx.ntt_dev(x.data_ptr(), m, False, True, 1); e1.record(ext); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    out[logn] = round(best, 3)
print(json.dumps(out))
