#!/bin/bash
mkdir -p gpurun_out
python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02b_bench_train_tf32.json 2> gpurun_out/r02b_train.err; echo "train rc=$?"
XATTN_ATTN_SIMT=1 python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02b_bench_train_tf32_simt_attn.json 2>> gpurun_out/r02b_train.err; echo "train simt rc=$?"
python scripts/prof_train.py 1024 3 tf32 > gpurun_out/prof_train_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02b_launches_train_tf32.csv python scripts/prof_train.py 1024 3 tf32 > gpurun_out/prof_train_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import json
for f in ("r02b_bench_train_tf32", "r02b_bench_train_tf32_simt_attn"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["ms_per_step"], d["value"], [(s["phase"], round(s["us"])) for s in d["roofline"]["stages"]])
PY
python scripts/launch_summary.py gpurun_out/r02b_launches_train_tf32.csv | sort -k1 -n -r | head -40
