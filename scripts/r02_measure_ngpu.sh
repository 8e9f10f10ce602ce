#!/bin/bash
# round 2: bench lines on N GPUs of one box (one rank per GPU):  bash scripts/r02_measure_ngpu.sh N
N=${1:-2}
mkdir -p gpurun_out
run() { timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
run 29511 --steps 30 --warmup 5 > gpurun_out/r02_bench_forward_${N}gpu.json 2> gpurun_out/r02_ngpu.err; echo "forward rc=$?"
run 29512 --workload train --steps 30 --warmup 5 > gpurun_out/r02_bench_train_tf32_${N}gpu.json 2>> gpurun_out/r02_ngpu.err; echo "train rc=$?"
run 29513 --workload train --no-overlap --steps 30 --warmup 5 > gpurun_out/r02_bench_train_tf32_${N}gpu_single_bucket.json 2>> gpurun_out/r02_ngpu.err; echo "train single bucket rc=$?"
run 29514 --workload sweep > gpurun_out/r02_bench_sweep_${N}gpu.json 2>> gpurun_out/r02_ngpu.err; echo "sweep rc=$?"
python -m pytest tests/test_gpu_multi.py -m gpu -q 2>&1 | tail -2
python - <<PY
import json
for f in ("forward", "train_tf32", "train_tf32", "sweep"):
    pass
for f in ("r02_bench_forward_${N}gpu", "r02_bench_train_tf32_${N}gpu", "r02_bench_train_tf32_${N}gpu_single_bucket", "r02_bench_sweep_${N}gpu"):
    try:
        d = json.loads([l for l in open(f"gpurun_out/{f}.json").read().splitlines() if l.startswith("{")][-1])
        print(f, "value", round(d["value"]), "ms", round(d["ms_per_step"], 4), "e2e", d.get("e2e") and round(d["e2e"]["value"]))
    except Exception as e:
        print(f, "FAILED", e)
PY
tail -c 800 gpurun_out/r02_ngpu.err
