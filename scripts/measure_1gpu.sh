# One-GPU measurement pass (run on the B200 box from the repo root: gpurun -- 'bash scripts/measure_1gpu.sh').
# Writes gpurun_out/m_*; the files judged are copied from there into profiles/ by hand (see profiles/README.md).
set -x
timeout -k 5 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout -k 5 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout -k 5 300 python bench.py --steps 50 --warmup 10 > gpurun_out/m_bench.json 2> gpurun_out/m_bench.err; tail -c 400 gpurun_out/m_bench.json
timeout -k 5 300 python bench.py --workload train --steps 20 --warmup 5 > gpurun_out/m_bench_train_tf32.json 2>> gpurun_out/m_bench.err; cut -c1-300 gpurun_out/m_bench_train_tf32.json
timeout -k 5 300 python bench.py --workload train --precision fp32 --steps 20 --warmup 5 > gpurun_out/m_bench_train_fp32.json 2>> gpurun_out/m_bench.err; cut -c1-300 gpurun_out/m_bench_train_fp32.json
timeout -k 5 300 python scripts/bench_configs.py sweep train flat > gpurun_out/m_configs_1gpu.jsonl 2>> gpurun_out/m_bench.err; wc -l gpurun_out/m_configs_1gpu.jsonl
# launch lists (each only after the same command exited 0 without ncu above)
timeout -k 5 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemm_tc|_tc_kernel|video|head|prior" -c 80 --csv --log-file gpurun_out/m_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/m_ncu.log 2>&1
tail -1 gpurun_out/m_ncu.log | cut -c1-200
timeout -k 5 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/m_launches_train_tf32.csv python scripts/prof_train.py 1024 3 tf32 > gpurun_out/m_ncu_train.log 2>&1
tail -1 gpurun_out/m_ncu_train.log
