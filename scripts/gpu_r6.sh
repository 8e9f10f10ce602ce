set -x
timeout 900 python -m pytest tests/test_gpu_bf16.py -q -x 2>&1 | tail -15
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r6_bench.json 2> gpurun_out/r6_bench.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r6_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'])
for s in d['roofline']['stages']: print(s['kernel'], round(s['us'],1), round(s['hbm_frac'],3))
PY
