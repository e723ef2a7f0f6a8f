set -e
python -m pytest tests -q -m gpu -x 2>&1 | tail -3
python bench.py > gpurun_out/r01d_bench_n1.json 2> gpurun_out/r01d_bench_n1.err
python bench.py --no-extras --steps 2 --warmup 1 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r01d_launches_prove_2p22.csv python bench.py --no-extras --steps 2 --warmup 1 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_accumulate|k_ntt_pass" -s 20 -c 8 -o gpurun_out/r01d_prof python bench.py --no-extras --steps 2 --warmup 1 > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
