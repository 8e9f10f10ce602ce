timeout -k 5 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
timeout -k 5 300 python scripts/bench_configs.py flat 2>&1 | cut -c1-600
