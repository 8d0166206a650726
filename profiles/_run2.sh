#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
S=gpurun_out/summary.txt
: > $S
timeout -k 10 600 python bench.py --steps 2000 --warmup 3 --no-cpu --no-sweep --no-cascade > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?" >> $S
python - >> $S <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print('value', d['value'], 'ms/step', d['ms_per_step'])
print('u8 train', d['paths']['u8_rows']['train_cache_off']['train_secs'], d['paths']['u8_rows']['train_cache_on']['train_secs'])
print('f32', d['paths']['f32_rows']['us_per_iter'], 'f64', d['paths']['f64_rows']['us_per_iter'])
print('parity u8', d['parity'].get('u8_rows',{}).get('meets_north_star'), d['parity'].get('u8_rows',{}).get('iterations'))
PY
timeout -k 10 300 python profiles/timeline_r2.py 60000 600 0 u8 > gpurun_out/timeline_u8.txt 2>&1; echo "timeline rc=$?" >> $S; head -19 gpurun_out/timeline_u8.txt >> $S
timeout -k 10 300 python profiles/timeline_r2.py 60000 3000 -1 u8 > gpurun_out/timeline_u8_cache.txt 2>&1; echo "timeline cache rc=$?" >> $S; grep -A15 "cached" gpurun_out/timeline_u8_cache.txt | head -17 >> $S
cat $S
