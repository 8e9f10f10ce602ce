set -x
timeout 600 python -m pytest tests/test_gpu_bf16.py -q -s -k "bf16_forward_config1 or ragged" 2>&1 | grep -v "^$" | tail -40
