set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1_pytest_gpu.log; tail -5 gpurun_out/r1_pytest_gpu.log
python __graft_entry__.py smoke 2>&1 | tail -3
python bench.py --precision fp32 --steps 10 --warmup 3 > gpurun_out/r1_bench_fp32.json 2> gpurun_out/r1_bench_fp32.err; tail -c 3000 gpurun_out/r1_bench_fp32.json; tail -5 gpurun_out/r1_bench_fp32.err
python bench.py --precision fp32 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r1_launches_fp32.csv python bench.py --precision fp32 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
