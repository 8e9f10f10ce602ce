# Two-GPU measurement pass (gpurun --gpus 2 -- 'bash scripts/measure_2gpu.sh'): one rank per GPU over NCCL.
set -x
timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 10 > gpurun_out/m_bench_2gpu.json 2> gpurun_out/m2.err; cut -c1-260 gpurun_out/m_bench_2gpu.json
timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload train --steps 20 --warmup 5 > gpurun_out/m_bench_train_2gpu.json 2>> gpurun_out/m2.err; cut -c1-260 gpurun_out/m_bench_train_2gpu.json
timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/bench_configs.py sweep train 2>> gpurun_out/m2.err | grep '^{' > gpurun_out/m_configs_2gpu.jsonl; wc -l gpurun_out/m_configs_2gpu.jsonl; tail -2 gpurun_out/m_configs_2gpu.jsonl | cut -c1-330
tail -3 gpurun_out/m2.err
