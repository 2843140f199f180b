timeout 900 python -m pytest tests/test_gpu_parser.py -m gpu -q -x 2>&1 | tail -25
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -8
