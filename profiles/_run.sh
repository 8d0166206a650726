timeout 200 python -m pytest tests -m gpu -x -q --tb=short --timeout 60 2>&1 | tail -5
echo "--- alone, fresh process"
timeout 100 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --tb=short --timeout 60 2>&1 | tail -5
