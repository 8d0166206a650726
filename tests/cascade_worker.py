"""One rank of a Cascade SVM run (launched by tests/test_cascade.py under torch.distributed.run).

  cascade_worker.py MODE N M OUT SOLVER     MODE = star | tree ; SOLVER = oracle (CPU checker as the per-rank
  solver: validates the merge protocol against the reference) | gpu (the product: C ABI on cuda:0)
Same data distribution as the reference (code/mpi_svm_main2.cpp:346-423): rank 0's view of all rows gives the
global min/max, rows are dealt in contiguous chunks of ceil(N/P), global id = row index."""
import json
import os
import sys

import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

mode, n, m, out_path, solver_kind = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5]
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
pkg = entry.load_package()
O = entry.load_oracle()          # the checker: scaling reference + (SOLVER=oracle) the per-rank solver
from svmb200 import parallel  # noqa: E402

X, dg = pkg.datagen.make_dataset(n)
y = pkg.datagen.binary_labels(dg)
Xt, dgt = pkg.datagen.make_dataset(m, split=1)
yt = pkg.datagen.binary_labels(dgt)

if solver_kind == "gpu":
    s = pkg.Solver(0)
    mn, mx = s.minmax_f64(X.astype(np.float64))                     # global statistics (mpi_svm_main2.cpp:413-419)
    start, end = parallel.partition(n, world, rank)
    X_part = s.scale_f64(X[start:end].astype(np.float64), mn, mx) if end > start else np.zeros((0, 784))
    solve = parallel.gpu_solve_fn(s, cache_rows=-1)
else:
    xs, mn, mx = O.scale(X)
    start, end = parallel.partition(n, world, rank)
    X_part = xs[start:end]

    def solve(Xa, ya, a0):
        r = O.smo_train(Xa, ya, alpha0=a0)
        return r["alpha"], r["b"]

ids = np.arange(start, end, dtype=np.int32)
fn = parallel.cascade_star if mode == "star" else parallel.cascade_tree
res = fn(X_part, y[start:end], ids, solve)
if rank == 0:
    Xs, ys, als, idsv = res["sv"]
    xt, _, _ = O.scale(Xt, mn, mx)
    # the reference's final prediction (mpi_svm_main2.cpp:721-733): s = -b + sum alpha_k y_k K(x, sv_k), pred = s >= 0
    if solver_kind == "gpu":
        s.set_train_scaled(Xs, ys)
        s.set_model(als, res["b"])
        correct, _ = s.predict_scaled(xt, yt, ge_zero=True)
    else:
        correct, _ = O.predict(xt, yt, Xs, ys, als, res["b"], ge_zero=True, tol=-1.0)
    json.dump({"b": res["b"], "sv_ids": sorted(int(i) for i in idsv), "rounds": res["rounds"], "converged": res["converged"],
               "correct": int(correct), "m": m}, open(out_path, "w"))
dist.barrier()
dist.destroy_process_group()
