for v in 0 3 4; do
  G16_ACC_VARIANT_G1=$v python bench.py --curve bls12_377 --log-n 22 --no-extras --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.readline());print('BLS VARIANT $v', round(d['value'],2), [round(x,2) for x in d['stages_ms']['accumulate']])"
done
