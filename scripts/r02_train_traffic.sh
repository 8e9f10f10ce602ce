#!/bin/bash
# DRAM bytes per launch of one Stage-1 step (B=1024, tf32): the plain run first, then ONE ncu metrics pass over the same command
mkdir -p gpurun_out
timeout 80 python scripts/prof_train.py 1024 4 tf32 > gpurun_out/prof_train_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_train_plain.log; exit 1; }
timeout 110 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/r02_traffic_train_tf32.csv python scripts/prof_train.py 1024 4 tf32 > gpurun_out/prof_train_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/prof_train_ncu.log; wc -l gpurun_out/r02_traffic_train_tf32.csv
