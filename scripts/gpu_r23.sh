timeout -k 5 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout -k 5 300 python scripts/bench_configs.py train 2>&1 | cut -c1-260
timeout -k 5 120 python scripts/prof_train.py 1024 3 tf32 && timeout -k 5 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r23_launches_train_tf32.csv python scripts/prof_train.py 1024 3 tf32 > gpurun_out/ncu_train.log 2>&1
tail -1 gpurun_out/ncu_train.log
