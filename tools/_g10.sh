timeout 900 python -m pytest tests/test_gpu_dropin.py -m gpu -q -x 2>&1 | grep -E "^E  |assert" | head -20
