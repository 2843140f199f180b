timeout 900 python -m pytest tests/test_gpu_dropin.py tests/test_gpu_golden_big.py -m gpu -q 2>&1 | tail -25
