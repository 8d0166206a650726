"""BASELINE.json configs[4]: modified two-layer (star) vs classical tree Cascade SVM across the GPUs of one box.
Launch: python -m torch.distributed.run --nproc-per-node P profiles/cascade_bench.py N
Each rank's sub-problems run on its own B200 through the C ABI; SV tuples move by NCCL all_gather (star) or
send/recv + broadcast (tree).  Rank 0 prints one JSON line per mode."""
import json, os, sys, time
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry

n = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
pkg = entry.load_package()
from svmb200 import parallel

X, dg = pkg.datagen.make_dataset(n); y = pkg.datagen.binary_labels(dg)
Xt, dgt = pkg.datagen.make_dataset(10000, split=1); yt = pkg.datagen.binary_labels(dgt)
s = pkg.Solver(local)
mn, mx = s.minmax_f64(X.astype(np.float64))
start, end = parallel.partition(n, world, rank)
X_part = s.scale_f64(X[start:end].astype(np.float64), mn, mx)
ids = np.arange(start, end, dtype=np.int32)
solve = parallel.gpu_solve_fn(s, cache_rows=-1)
# the undivided problem on one GPU, for comparison
s.set_train_u8(X, y)
t0 = time.perf_counter(); full = s.smo_train(cache_rows=-1); t_full = time.perf_counter() - t0
sv_full = set(s.sv_ids().tolist())
for mode, fn in (("star", parallel.cascade_star), ("tree", parallel.cascade_tree)):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = fn(X_part, y[start:end], ids, solve)
    dist.barrier(); torch.cuda.synchronize()
    secs = time.perf_counter() - t0
    if rank == 0:
        Xs, ys, als, idsv = res["sv"]
        s.set_train_scaled(Xs, ys); s.set_model(als, res["b"])
        xt = s.scale_f64(Xt.astype(np.float64), mn, mx)
        correct, _ = s.predict_scaled(xt, yt, ge_zero=True)
        sv = set(int(i) for i in idsv)
        print(json.dumps({"mode": mode, "ranks": world, "n": n, "secs": round(secs, 3), "rounds": res["rounds"], "converged": res["converged"],
                          "n_sv": len(sv), "b": res["b"], "test_correct_of_10000": int(correct),
                          "undivided_1gpu": {"secs": round(t_full, 3), "n_sv": len(sv_full), "b": full["b"], "iterations": full["num_iter"] - 1},
                          "sv_set_equals_undivided": sv == sv_full, "sv_symmetric_difference": len(sv ^ sv_full)}), flush=True)
dist.destroy_process_group()
