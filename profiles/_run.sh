timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --tb=short --timeout 100 2>&1 | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 profiles/cascade_bench.py 60000 2> gpurun_out/cascade8.err | grep -v NCCL | tee gpurun_out/cascade_r1_8gpu.jsonl
tail -3 gpurun_out/cascade8.err | grep -v OMP
