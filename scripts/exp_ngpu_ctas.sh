#!/bin/bash
# Data-parallel Stage-1 step with the persistent kernels' grids capped (XATTN_MAX_CTAS), leaving SMs to the NCCL kernels that run
# beside the input-projection backward.  usage: exp_ngpu_ctas.sh N "148 140 132"
N=${1:-2}
mkdir -p gpurun_out
for cap in ${2:-148 140 132}; do
  XATTN_MAX_CTAS=$cap timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --workload train --steps 30 --warmup 5 --no-cpu-baseline 2>gpurun_out/ctas.err | tail -1 > gpurun_out/ctas_${N}gpu_$cap.json
  python -c "
import json; d=json.loads(open('gpurun_out/ctas_${N}gpu_$cap.json').read()); print('cap', $cap, d['ms_per_step'], d['value'])"
done
