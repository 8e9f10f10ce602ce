# N-GPU bench lines (gpurun --gpus N -- 'bash scripts/measure_ngpu.sh N'): forward and Stage-1 step, one rank per GPU.
set -x
N=${1:-8}
timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 50 --warmup 10 2> gpurun_out/m${N}.err | grep '^{' > gpurun_out/m_bench_${N}gpu.json; cut -c1-260 gpurun_out/m_bench_${N}gpu.json
timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload train --steps 20 --warmup 5 2>> gpurun_out/m${N}.err | grep '^{' > gpurun_out/m_bench_train_${N}gpu.json; cut -c1-260 gpurun_out/m_bench_train_${N}gpu.json
python - <<PY
import json
d=json.loads(open("gpurun_out/m_bench_${N}gpu.json").read().strip().splitlines()[-1]); print("e2e", d["e2e"])
PY
for i in $(seq 0 $((N-1))); do a=$(nvidia-smi -i $i --query-gpu=pci.bus_id --format=csv,noheader | tr 'A-Z' 'a-z' | sed 's/^0000//'); echo "gpu $i $a numa $(cat /sys/bus/pci/devices/$a/numa_node 2>&1)"; done
tail -3 gpurun_out/m${N}.err
