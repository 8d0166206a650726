#!/bin/bash
# Multi-GPU part of the evidence (gpurun --gpus N -- bash profiles/collect_multi_gpu.sh): the sharding tests on an 8-GPU box, then
# the bench as the driver launches it (--steps 20) and with --steps 2000.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
S=gpurun_out/summary_mg$NG.txt
: > $S
echo "gpus $NG" >> $S
if [ $NG = 8 ]; then
  timeout -k 10 900 python -m pytest tests/test_multi_gpu.py -m gpu -k bit_identical -x -q > gpurun_out/pytest_mg$NG.log 2>&1; echo "pytest rc=$?" >> $S; tail -3 gpurun_out/pytest_mg$NG.log >> $S
  EXTRA=""
else
  EXTRA="--no-cpu"
fi
run() { K=$1; shift
  timeout -k 10 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $((29600+NG+K%7)) bench.py --gpus $NG --steps $K --warmup 3 "$@" > gpurun_out/bench_r2_n${NG}_steps$K.json 2> gpurun_out/bench_n${NG}_$K.err
  echo "bench n=$NG steps=$K rc=$?" >> $S
}
run 20 $EXTRA
run 2000 --no-cascade --no-cpu
cat $S
