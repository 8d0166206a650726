"""One rank of a row-sharded SMO run (launched by tests/test_multi_gpu.py / bench under torch.distributed.run).
Each rank owns a GPU; X is uploaded in full everywhere; rows are split contiguously; the per-iteration
exchange runs inside the persistent kernels over NVLink peer memory."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

n, out_path = int(sys.argv[1]), sys.argv[2]
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
pkg = entry.load_package()
from svmb200 import parallel  # noqa: E402

X, dg = pkg.datagen.make_dataset(n)
y = pkg.datagen.binary_labels(dg)
res = {}
# single-GPU run of the same problem on this rank's GPU: the sharded result must equal it bit for bit
single = pkg.Solver(local)
single.set_train_u8(X, y)
ref = single.smo_train(trace_cap=100000)
ref_trace = single.get_trace(100000)
ref_cached = single.smo_train(cache_rows=-1)
single.close()

s = pkg.Solver(local)
s.set_train_u8(X, y)
sh = parallel.ShardedSolver(s)
for label, kw in (("cache_off", dict(cache_rows=0)), ("cache_on", dict(cache_rows=-1)), ("cache_small", dict(cache_rows=32))):
    out = sh.smo_train(trace_cap=100000, **kw)
    res[label] = {"alpha_equal": bool(np.array_equal(out["alpha"], ref["alpha"])), "b_equal": out["b"] == ref["b"],
                  "iters": out["num_iter"], "ref_iters": ref["num_iter"], "stop": out["stop_reason"],
                  "trace_equal": bool(np.array_equal(s.get_trace(100000), ref_trace)) if rank == 0 else True,
                  "x_passes": out["x_passes"], "cache_hits": out["cache_hits"], "us_per_iter": 1e3 * out["smo_ms"] / max(out["num_iter"] - 1, 1),
                  "sv_equal": bool(np.array_equal(s.sv_ids(), np.nonzero(ref["alpha"] > 1e-8)[0]))}
Xt, dgt = pkg.datagen.make_dataset(200, split=1)
res["correct"] = s.predict_u8(Xt, pkg.datagen.binary_labels(dgt))[0]
allres = [None] * world
dist.all_gather_object(allres, res)
if rank == 0:
    json.dump(allres, open(out_path, "w"))
dist.barrier()
dist.destroy_process_group()
