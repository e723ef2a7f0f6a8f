python -m pytest tests -q -m gpu 2>&1 | tail -3
python __graft_entry__.py smoke 2>&1 | tail -2
python bench.py > gpurun_out/r01e_bench_n1.json 2> gpurun_out/r01e_bench_n1.err; tail -c 300 gpurun_out/r01e_bench_n1.err
python -c "
import json;d=json.load(open('gpurun_out/r01e_bench_n1.json'));print(d['value'],d['e2e'],d['roofline']['frac'],d['gpu_launches'],d['clocks'])"
python bench.py --impl reference --steps 1 --warmup 0 2>/dev/null | tail -1 | cut -c1-400
