set -x
python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^$" | tail -15 > gpurun_out/r2_pytest_gpu.log; tail -12 gpurun_out/r2_pytest_gpu.log
python __graft_entry__.py smoke 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_bf16.json 2> gpurun_out/r2_bench_bf16.err; tail -c 2500 gpurun_out/r2_bench_bf16.json; tail -5 gpurun_out/r2_bench_bf16.err
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r2_launches_bf16.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
