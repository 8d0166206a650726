"""Cascade SVM across the GPUs of one box -- the program that stands where the reference's MPI programs stand.

  python -m torch.distributed.run --nproc-per-node P --master-addr 127.0.0.1 \
      <package dir>/cascade_main.py --mode star --train-csv mnist3_train_data.csv --test-csv mnist3_test_data.csv

Follows `main()` of code/mpi_svm_main2.cpp (star, :301-780) and code/mpi_svm_main3.cpp (tree, :403-838): rank 0 reads
the data (here every rank reads it: one box, one file system), min/max + scaling on the training set, contiguous
partitions with the row index as global id (mpi_svm_main2.cpp:346-402), Cascade rounds until the global SV-ID set
repeats (at most 50), the model evaluated on the test set with the ">= 0" rule (:721-733), and the reference's report
lines.  One process per GPU; every sub-problem is solved by the persistent SMO kernel through the C ABI; SV tuples move
over torch.distributed (NCCL).  Inputs: the reference's CSV (header line, last column = digit, label = digit == 1,
main3.cpp:13-54), MNIST idx files, or --synthetic N (the generator bench.py uses)."""
import argparse
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
import __graft_entry__ as entry  # noqa: E402


def read_reference_csv(path):
    """code/main3.cpp:13-54: first line discarded, every other line = features..., digit."""
    data = np.loadtxt(path, delimiter=",", skiprows=1, dtype=np.float64, ndmin=2)
    return np.ascontiguousarray(data[:, :-1]), data[:, -1].astype(np.int64)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--mode", choices=("star", "tree"), default="star", help="star = modified two-layer cascade (mpi_svm_main2.cpp), tree = classical (mpi_svm_main3.cpp)")
    ap.add_argument("--train-csv"), ap.add_argument("--test-csv")
    ap.add_argument("--train-images"), ap.add_argument("--train-labels"), ap.add_argument("--test-images"), ap.add_argument("--test-labels")
    ap.add_argument("--synthetic", type=int, default=0, metavar="N", help="N synthetic MNIST-shaped training rows + 10 000 test rows")
    ap.add_argument("--digit", type=int, default=1, help="the positive class (the reference: 1)")
    ap.add_argument("--C", type=float, default=10.0), ap.add_argument("--gamma", type=float, default=0.00125)
    ap.add_argument("--max-rounds", type=int, default=50)
    args = ap.parse_args(argv)

    if not torch.cuda.is_available():
        sys.exit("cascade_main.py needs CUDA devices: one process per GPU")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    if args.mode == "tree" and world & (world - 1):
        sys.exit("the tree cascade needs a power-of-two number of ranks (mpi_svm_main3.cpp:420-428)")
    pkg = entry.load_package()
    from svmb200 import parallel

    def say(msg):
        if rank == 0:
            print(f"[rank 0] {msg}", flush=True)

    if args.synthetic:
        X, dg = pkg.datagen.make_dataset(args.synthetic)
        Xt, dgt = pkg.datagen.make_dataset(10000, split=1)
    elif args.train_csv:
        X, dg = read_reference_csv(args.train_csv)
        Xt, dgt = read_reference_csv(args.test_csv)
    elif args.train_images:
        X, dg = pkg.datagen.read_idx(args.train_images, args.train_labels)
        Xt, dgt = pkg.datagen.read_idx(args.test_images, args.test_labels)
    else:
        ap.error("give --train-csv/--test-csv, --train-images/... or --synthetic N")
    y = pkg.datagen.binary_labels(np.asarray(dg), args.digit)
    yt = pkg.datagen.binary_labels(np.asarray(dgt), args.digit)
    n, d = X.shape
    say(f"n = {n}")
    say(f"n_features = {d}")

    t_start = time.perf_counter()
    s = pkg.Solver(local)
    X64 = np.ascontiguousarray(X, dtype=np.float64)
    mn, mx = s.minmax_f64(X64)                                                  # find_min_max on the training set
    start, end = parallel.partition(n, world, rank)
    X_part = s.scale_f64(X64[start:end], mn, mx)                                # scale_features with the training statistics
    ids = np.arange(start, end, dtype=np.int32)
    solve = parallel.gpu_solve_fn(s, C=args.C, gamma=args.gamma, cache_rows=-1)
    fn = parallel.cascade_star if args.mode == "star" else parallel.cascade_tree
    res = fn(X_part, y[start:end], ids, solve, max_rounds=args.max_rounds, log=say)
    dist.barrier()
    torch.cuda.synchronize()
    t_train = time.perf_counter()

    Xs, ys, als, idsv = res["sv"]
    if res["converged"]:
        # the star program counts rounds from 0 (mpi_svm_main2.cpp:436,681,768), the tree program from 1 (mpi_svm_main3.cpp:562,568,750)
        say(f"Converged at round {res['rounds'] - 1 if args.mode == 'star' else res['rounds']}, SV count = {len(idsv)}")
    else:
        say(f"Not converged after {res['rounds']} rounds, SV count = {len(idsv)}")
    say(f"Final b = {res['b']:.15f}")
    correct = 0
    if rank == 0 and len(idsv) == 0:
        say("Final model has no support vectors: nothing to predict with")
    elif rank == 0:
        s.set_train_scaled(Xs, ys)
        s.set_model(als, res["b"])
        xt = s.scale_f64(np.ascontiguousarray(Xt, dtype=np.float64), mn, mx)
        correct, _ = s.predict_scaled(xt, yt, ge_zero=True)                     # ">= 0" as the MPI programs (mpi_svm_main2.cpp:729)
        say(f"Test accuracy (final model) = {correct / len(yt):.15f} ({correct}/{len(yt)})")
        say(f"Final global SV count = {len(idsv)}")
        say(f"Cascade finished in {res['rounds']} rounds")
    t_end = time.perf_counter()
    say(f"training time = {1e3 * (t_train - t_start):.0f} ms")
    say(f"prediction time = {1e3 * (t_end - t_train):.0f} ms")
    say(f"elapsed time = {1e3 * (t_end - t_start):.0f} ms")
    s.close()
    dist.barrier()
    dist.destroy_process_group()
    return {"n_sv": len(idsv), "b": res["b"], "rounds": res["rounds"], "correct": int(correct)}


if __name__ == "__main__":
    main()
