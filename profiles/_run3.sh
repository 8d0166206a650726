#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
S=gpurun_out/summary_mg.txt
: > $S
NG=$(nvidia-smi -L | wc -l); echo "gpus $NG" >> $S
timeout -k 10 900 python -m pytest tests/test_multi_gpu.py tests/test_cascade_nccl.py -m gpu -x -q > gpurun_out/pytest_mg.log 2>&1; echo "pytest mg rc=$?" >> $S; tail -4 gpurun_out/pytest_mg.log >> $S
for K in 20 2000; do
timeout -k 10 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $((29700+K%7)) bench.py --gpus $NG --steps $K --warmup 3 --no-cpu > gpurun_out/bench_mg_n${NG}_steps$K.json 2> gpurun_out/bench_mg_$K.err; echo "bench n=$NG steps=$K rc=$?" >> $S
python - $NG $K >> $S <<'PY'
import json,sys
d=json.load(open(f'gpurun_out/bench_mg_n{sys.argv[1]}_steps{sys.argv[2]}.json'))
print('value', d['value'], 'ms/step', d['ms_per_step'], d.get('stop_reason'))
print('mg', json.dumps(d.get('multi_gpu'))[:700])
c=d.get('cascade') or {}
print('cascade', {k:(c[k]['secs'],c[k]['rounds'],c[k]['n_sv'],c[k]['sv_set_equals_undivided_reference']) for k in ('star','tree') if k in c})
print('train', d['train']['cache_off']['train_secs'], d['train']['cache_on']['train_secs'])
PY
done
cat $S
