set -x
timeout -k 5 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
timeout -k 5 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout -k 5 300 python bench.py --steps 50 --warmup 10 > gpurun_out/r10_bench.json 2> gpurun_out/r10_bench.err; tail -c 600 gpurun_out/r10_bench.json
timeout -k 5 120 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && timeout -k 5 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemm_tc|_tc_kernel|video|head|prior" -c 80 --csv --log-file gpurun_out/r10_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
