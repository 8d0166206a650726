timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --tb=short --timeout 200 2>&1 | tail -12
