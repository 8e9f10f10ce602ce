set -x
timeout -k 5 600 python scripts/bench_configs.py sweep train flat > gpurun_out/r16_configs_1gpu.jsonl 2> gpurun_out/r16_configs.err; tail -3 gpurun_out/r16_configs.err; cat gpurun_out/r16_configs_1gpu.jsonl | cut -c1-420
timeout -k 5 120 python scripts/prof_train.py 1024 3 && timeout -k 5 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r16_launches_train.csv python scripts/prof_train.py 1024 3 > gpurun_out/ncu_train.log 2>&1
tail -2 gpurun_out/ncu_train.log
