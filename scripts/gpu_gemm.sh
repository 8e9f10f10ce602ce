set -x
timeout 300 python -m pytest tests/test_gpu_bf16.py -x -q -k gemm 2>&1 | tail -25
