timeout -k 5 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
timeout -k 5 300 python bench.py --steps 50 --warmup 10 --no-cpu-baseline > gpurun_out/r20_bench.json 2> gpurun_out/r20_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r20_bench.json').read().strip().splitlines()[-1]); print('PDL   ', d['value'], d['ms_per_step'])"
XATTN_NO_PDL=1 timeout -k 5 300 python bench.py --steps 50 --warmup 10 --no-cpu-baseline > gpurun_out/r20_bench_nopdl.json 2>> gpurun_out/r20_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r20_bench_nopdl.json').read().strip().splitlines()[-1]); print('no PDL', d['value'], d['ms_per_step'])"
timeout -k 5 300 python scripts/bench_configs.py sweep 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(d['B_total'], round(d['ms_per_step']*1e3,1), d['ms_per_step_cuda_graph'] and round(d['ms_per_step_cuda_graph']*1e3,1))"
