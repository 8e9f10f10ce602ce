set -x
python scripts/prof_fwd.py 4096 3 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"head_pooled|prior_kernel|video_update|video_pre" -s 8 -c 4 -o gpurun_out/r6_prof python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
