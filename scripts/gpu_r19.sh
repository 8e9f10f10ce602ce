timeout -k 5 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -8
timeout -k 5 120 python scripts/probe_latency.py 2>&1 | grep "^B="
timeout -k 5 300 python scripts/bench_configs.py sweep 2>&1 | cut -c1-330
