#!/bin/bash
# ncu --set full capture of the four mma.sync attention kernels of one Stage-1 step (B=1024, tf32), with source correlation
mkdir -p gpurun_out
python scripts/prof_train.py 1024 3 tf32 > gpurun_out/prof_train_plain.log 2>&1 && \
timeout 600 ncu --set full --import-source on --clock-control none -k regex:attn_.*mma --launch-skip 8 -c 4 -f -o gpurun_out/r02_attn_mma \
  python scripts/prof_train.py 1024 3 tf32 > gpurun_out/attn_ncu.log 2>&1; echo "ncu rc=$?"
tail -5 gpurun_out/attn_ncu.log
ls -la gpurun_out/r02_attn_mma.ncu-rep
