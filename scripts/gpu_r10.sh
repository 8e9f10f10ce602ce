timeout -k 5 200 python -m pytest tests/test_gpu_bf16.py -q -x -s -k "full_batch or config1 or ragged" 2>&1 | grep -v "^$" | tail -25
