set -x
timeout -k 5 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
timeout -k 5 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout -k 5 300 python bench.py --steps 50 --warmup 10 > gpurun_out/r15_bench.json 2> gpurun_out/r15_bench.err; tail -c 3000 gpurun_out/r15_bench.json
timeout -k 5 120 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && timeout -k 5 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemm_tc|_tc_kernel|video|head|prior" -c 80 --csv --log-file gpurun_out/r15_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
timeout -k 5 120 python scripts/prof_fwd.py 4096 3 && timeout -k 5 400 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_kernel|_tc_kernel|video|head|prior" -s 16 -c 8 -o gpurun_out/r15_full -f python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
