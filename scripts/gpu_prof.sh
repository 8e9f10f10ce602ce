set -x
timeout -k 5 120 python scripts/prof_fwd.py 4096 3 > gpurun_out/plain.log 2>&1 && timeout -k 5 300 ncu --set full --clock-control none --import-source on -k regex:"v2a_tc" -s 2 -c 1 -o gpurun_out/r9_prof python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
