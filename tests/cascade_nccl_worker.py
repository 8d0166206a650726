"""One rank of a device-resident Cascade SVM run through the C ABI (svmb200_cascade_*, NCCL between the GPUs), launched by
tests/test_cascade_nccl.py under torch.distributed.run (gloo: it only carries the 128-byte NCCL id).

  cascade_nccl_worker.py MODE N M OUT"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

mode, n, m, out_path = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
pkg = entry.load_package()
from svmb200 import cascade_bench  # noqa: E402

X, dg = pkg.datagen.make_dataset(n)
y = pkg.datagen.binary_labels(dg)
Xt, dgt = pkg.datagen.make_dataset(m, split=1)
yt = pkg.datagen.binary_labels(dgt)
s = cascade_bench.connect(pkg, local, rank, world)
out = cascade_bench.run_gpu(s, X, y, rank, world, mode)
ids, alpha = s.cascade_get_model()
correct, _ = s.predict_u8(Xt, yt, ge_zero=True)          # every rank holds the model: ">= 0" as the MPI programs (mpi_svm_main2.cpp:729)
res = {"b": out["b"], "sv_ids": sorted(int(i) for i in ids), "rounds": out["rounds"], "converged": bool(out["converged"]),
       "correct": int(correct), "n_sv": out["n_sv"], "alpha_sum": float(alpha.sum()), "secs": out["secs"]}
allres = [None] * world
dist.all_gather_object(allres, res)
if rank == 0:
    json.dump(allres, open(out_path, "w"))
s.close()
dist.barrier()
dist.destroy_process_group()
