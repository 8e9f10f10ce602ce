timeout -k 5 120 python scripts/prof_fwd.py 4096 3 > gpurun_out/plain.log 2>&1 && timeout -k 5 600 ncu --set full --clock-control none --import-source on -k regex:"$1" -s ${2:-3} -c ${3:-3} -o gpurun_out/${4:-r14_prof} -f python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
