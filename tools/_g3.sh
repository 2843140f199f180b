timeout 300 python -m pytest tests/test_gpu_eigh_cluster.py -m gpu -x -q 2>&1 | tail -30
timeout 600 python -m pytest tests/test_gpu_golden_big.py -m gpu -q 2>&1 | tail -15
