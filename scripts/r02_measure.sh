#!/bin/bash
# round 2: bench lines of every workload + the ncu launch list (times and DRAM bytes) of the bf16 forward
mkdir -p gpurun_out
python bench.py --steps 30 --warmup 5 > gpurun_out/r02_bench_forward.json 2> gpurun_out/r02_bench_forward.err; echo "forward rc=$?"
python bench.py --workload train --steps 20 --warmup 5 > gpurun_out/r02_bench_train_tf32.json 2> gpurun_out/r02_bench_train.err; echo "train rc=$?"
python bench.py --workload train --precision fp32 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_train_fp32.json 2>> gpurun_out/r02_bench_train.err; echo "train fp32 rc=$?"
python bench.py --workload sweep > gpurun_out/r02_bench_sweep.json 2> gpurun_out/r02_bench_sweep.err; echo "sweep rc=$?"
python bench.py --workload flat --steps 30 --warmup 5 > gpurun_out/r02_bench_flat.json 2> gpurun_out/r02_bench_flat.err; echo "flat rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err; echo "ref rc=$?"
python scripts/prof_fwd.py 4096 3 > gpurun_out/prof_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"gemm_tc|_tc_kernel|video|head|prior" -c 60 --csv --log-file gpurun_out/r02_launches_bf16.csv python scripts/prof_fwd.py 4096 3 > gpurun_out/prof_ncu.log 2>&1; echo "ncu rc=$?"
tail -c 600 gpurun_out/*.err
cut -c1-600 gpurun_out/r02_bench_forward.json
