timeout -k 5 200 python -m pytest tests/test_gpu_bf16.py -q -x 2>&1 | tail -5
timeout -k 5 120 python scripts/prof_fwd.py 4096 3 > gpurun_out/plain.log 2>&1 && timeout -k 5 600 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_kernel" -s 6 -c 4 -o gpurun_out/r13_prof -f python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
