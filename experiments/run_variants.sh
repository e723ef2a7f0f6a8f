set -x
for v in "0 0" "3 0" "4 0" "0 3" "0 4"; do
  set -- $v
  G16_ACC_VARIANT_G1=$1 G16_ACC_VARIANT_G2=$2 python bench.py --no-extras --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.readline());print('VARIANT $1 $2', round(d['value'],2), [round(x,2) for x in d['stages_ms']['accumulate']])"
done
G16_ACC_VARIANT_G1=3 G16_ACC_VARIANT_G2=3 python -m pytest tests/test_gpu_msm.py tests/test_gpu_prove.py -q -m gpu -x 2>&1 | tail -3
