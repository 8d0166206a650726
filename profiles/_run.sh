timeout 150 python -m pytest tests -m gpu -q --tb=short --timeout 60 2>&1 | grep -E "passed|failed|FAILED|assert|Error" | head -20
for ks in 4 2; do echo "KSLOTS=$ks"; SVMB200_KSLOTS=$ks timeout 100 python profiles/sweep.py 60000,10000,512 4 0 1000; done 2>&1 | tee gpurun_out/sweep9.log
timeout 60 python profiles/timeline.py 512 340 2>&1 | tee gpurun_out/timeline_512.log
timeout 60 python profiles/timeline.py 60000 600 2>&1 | tee gpurun_out/timeline_60k.log
