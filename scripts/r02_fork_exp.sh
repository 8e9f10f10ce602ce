#!/bin/bash
# Stage-1 step with and without the side-stream overlap (prior kernels beside the v2a input-projection GEMMs)
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tf32.py tests/test_gpu_rng.py tests/test_zz_no_alloc.py -q -x > gpurun_out/fork_tests.log 2>&1; tail -3 gpurun_out/fork_tests.log
for i in 1 2; do
python bench.py --workload train --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/fork_on_$i.json 2> gpurun_out/fork.err; echo "on rc=$?"
XATTN_NO_FORK=1 python bench.py --workload train --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/fork_off_$i.json 2>> gpurun_out/fork.err; echo "off rc=$?"
done
python - <<'PY'
import json
for f in ("fork_on_1", "fork_off_1", "fork_on_2", "fork_off_2"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d["ms_per_step"], d["value"], [(s["phase"], round(s["us"])) for s in d["roofline"]["stages"]])
PY
tail -c 300 gpurun_out/fork.err
