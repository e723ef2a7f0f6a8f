"""prints the measured integer-pipe rates of this GPU (roofline denominators)."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gnark_icicle_b200 import lib
c = lib.Context(0)
imad, wide = c.imad_peak()
wx, frmul, fpmul = c.pipe_rates()
print(json.dumps({"imad_per_s": imad, "imad_wide_per_s": wide, "imad_wide_carry_per_s": wx, "fr_mul_per_s": frmul}))
