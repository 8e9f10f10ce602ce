set -x
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py --steps 30 --warmup 5 > gpurun_out/r5_bench.json 2> gpurun_out/r5_bench.err; tail -c 3000 gpurun_out/r5_bench.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemm_tc|attn|video|head" -c 60 --csv --log-file gpurun_out/r5_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
