set -x
timeout -k 5 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout -k 5 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout -k 5 300 python bench.py --steps 50 --warmup 10 > gpurun_out/r24_bench.json 2> gpurun_out/r24_bench.err; tail -c 400 gpurun_out/r24_bench.json
timeout -k 5 300 python bench.py --workload train --steps 20 --warmup 5 > gpurun_out/r24_bench_train_tf32.json 2>> gpurun_out/r24_bench.err; cut -c1-300 gpurun_out/r24_bench_train_tf32.json
timeout -k 5 300 python bench.py --workload train --precision fp32 --steps 20 --warmup 5 > gpurun_out/r24_bench_train_fp32.json 2>> gpurun_out/r24_bench.err; cut -c1-300 gpurun_out/r24_bench_train_fp32.json
for pdl in 0 1; do
  if [ $pdl = 0 ]; then export XATTN_NO_PDL=1; else unset XATTN_NO_PDL; fi
  timeout -k 5 300 python bench.py --batch 16384 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('B=16384 pdl=$pdl', d['value'], d['ms_per_step'])"
done
unset XATTN_NO_PDL
timeout -k 5 300 python scripts/bench_configs.py sweep train flat > gpurun_out/r24_configs_1gpu.jsonl 2>> gpurun_out/r24_bench.err; wc -l gpurun_out/r24_configs_1gpu.jsonl
timeout -k 5 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemm_tc|_tc_kernel|video|head|prior" -c 80 --csv --log-file gpurun_out/r24_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -1 gpurun_out/ncu.log | cut -c1-200
timeout -k 5 400 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_kernel|_tc_kernel|video|head|prior" -s 16 -c 8 -o gpurun_out/r24_full -f python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_full.log
timeout -k 5 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r24_launches_train_tf32.csv python scripts/prof_train.py 1024 3 tf32 > gpurun_out/ncu_train.log 2>&1
tail -1 gpurun_out/ncu_train.log
