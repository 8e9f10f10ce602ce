timeout 300 python -m pytest tests/test_gpu_bf16.py -q -k "umma_probe" 2>&1 | tail -30
