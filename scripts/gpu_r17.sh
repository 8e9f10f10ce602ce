timeout -k 5 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none -c 120 --csv --log-file gpurun_out/r17_launches_flat.csv python scripts/bench_configs.py flat > gpurun_out/ncu_flat.log 2>&1
tail -2 gpurun_out/ncu_flat.log
python - <<'PY'
import csv,re,collections
rows=list(csv.reader(open('gpurun_out/r17_launches_flat.csv')))
i=[k for k,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[i]; K=h.index('Kernel Name'); V=h.index('Metric Value'); N=h.index('Metric Name'); G=h.index('Grid Size')
for r in rows[i+1:i+1+40]:
    if r[N]=='gpu__time_duration.sum': print(r[V], r[G], r[K][:90])
PY
