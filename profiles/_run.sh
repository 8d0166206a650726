timeout 150 python -m pytest tests -m gpu -q --tb=short --timeout 60 -s 2>&1 | grep -E "passed|failed|FAILED|assert|prefix|Error" | head -20
timeout 100 python profiles/sweep.py 60000,10000,512 4,1 0,1 1000 2>&1 | tee gpurun_out/sweep6.log
timeout 60 python profiles/timeline.py 60000 600 2>&1 | tee gpurun_out/timeline_60k.log
timeout 60 python profiles/timeline.py 60000 3000 -1 2>&1 | tee gpurun_out/timeline_60k_cache.log
