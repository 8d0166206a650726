timeout 400 python -m pytest tests -m gpu -x -q --tb=short --timeout 200 2>&1 | tail -8
