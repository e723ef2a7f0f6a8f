for v in "0 0" "0 5" "0 6" "0 7"; do
  set -- $v
  G16_ACC_VARIANT_G1=$1 G16_ACC_VARIANT_G2=$2 python bench.py --no-extras --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.readline());print('VARIANT $1 $2', round(d['value'],2), [round(x,2) for x in d['stages_ms']['accumulate']])"
done
