set -x
timeout 600 python -m pytest tests/test_gpu_bf16.py -q -x 2>&1 | tail -5
python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>&1 | tail -c 1200
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemm_tc|attn|video|head" -c 60 --csv --log-file gpurun_out/r4_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
