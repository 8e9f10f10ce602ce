set -x
timeout -k 5 120 python scripts/prof_fwd.py 4096 3 > gpurun_out/plain.log 2>&1 && timeout -k 5 600 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_kernel|_tc_kernel|video_update|video_pre" -s 12 -c 6 -o gpurun_out/r10_prof python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
