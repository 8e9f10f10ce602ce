#!/bin/bash
# round 2, final single-GPU measurements: one bench line per BASELINE config + ncu launch lists (forward with DRAM bytes, train)
mkdir -p gpurun_out
python bench.py --steps 50 --warmup 10 > gpurun_out/r02_bench_forward.json 2> gpurun_out/r02_final.err; echo "forward rc=$?"
python bench.py --workload train --steps 30 --warmup 5 > gpurun_out/r02_bench_train_tf32.json 2>> gpurun_out/r02_final.err; echo "train rc=$?"
python bench.py --workload train --precision fp32 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_train_fp32.json 2>> gpurun_out/r02_final.err; echo "train fp32 rc=$?"
python bench.py --workload sweep > gpurun_out/r02_bench_sweep.json 2>> gpurun_out/r02_final.err; echo "sweep rc=$?"
python bench.py --workload flat --steps 30 --warmup 5 > gpurun_out/r02_bench_flat.json 2>> gpurun_out/r02_final.err; echo "flat rc=$?"
python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/r02_bench_reference.json 2>> gpurun_out/r02_final.err; echo "ref rc=$?"
python bench.py --impl reference --workload train --steps 5 --warmup 1 > gpurun_out/r02_bench_reference_train.json 2>> gpurun_out/r02_final.err; echo "ref train rc=$?"
python scripts/prof_train.py 1024 3 tf32 > gpurun_out/prof_train_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_train_tf32.csv python scripts/prof_train.py 1024 3 tf32 > gpurun_out/prof_train_ncu.log 2>&1; echo "ncu train rc=$?"
tail -c 500 gpurun_out/r02_final.err
