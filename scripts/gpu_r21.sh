timeout -k 5 300 python -m pytest tests/test_gpu_tf32.py -m gpu -q -s 2>&1 | grep -v "^E  \|^    \|^$" | tail -30
