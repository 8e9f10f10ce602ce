timeout -k 5 150 python -m pytest tests/test_gpu_bf16.py -q -x 2>&1 | tail -5
timeout -k 5 120 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r8_bench.json 2> gpurun_out/r8_bench.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r8_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['whole_step']['bracketed_pass_ms_per_step'])
print([(s['kernel'], round(s['us'],1)) for s in d['roofline']['stages']])
PY
bash scripts/gpu_ll.sh 2>&1 | grep "xf::"
