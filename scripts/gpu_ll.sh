# launch list (per-kernel durations) of the bf16 forward at B=4096
timeout -k 5 120 python scripts/prof_fwd.py 4096 3 > gpurun_out/plain.log 2>&1 && timeout -k 5 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ll.csv python scripts/prof_fwd.py 4096 3 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
python scripts/launch_summary.py gpurun_out/ll.csv
