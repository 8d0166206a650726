"""BASELINE.json configs[4] for bench.py: modified two-layer (star) vs classical tree Cascade SVM over the N GPUs of the
run, through the C ABI (svmb200_cascade_*: NCCL all-gather / send-recv / broadcast of the support-vector tuples, device
resident), and the reference's own MPI programs (code/mpi_svm_main2.cpp, mpi_svm_main3.cpp, unmodified, built against
oracle/mpi_shim) timed on the box's host cores on a bounded sample of the same rows.

torch.distributed only hands the 128-byte NCCL id from rank 0 to the others."""
from __future__ import annotations

import os
import re
import subprocess
import tempfile
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = {"star": os.path.join(ROOT, "oracle", "_ref", "mpi_svm2"), "tree": os.path.join(ROOT, "oracle", "_ref", "mpi_svm3")}


def partition(n_total: int, world: int, rank: int):
    """Contiguous chunks of ceil(n/P) rows; global id = original row index (mpi_svm_main2.cpp:346-402)."""
    chunk = (n_total + world - 1) // world
    start = min(rank * chunk, n_total)
    return start, max(start, min((rank + 1) * chunk, n_total))


def connect(pkg, local_rank: int, rank: int, world: int):
    """A context on this rank's GPU joined to the cascade communicator."""
    from .binding import cascade_unique_id
    box = [cascade_unique_id() if rank == 0 else None]
    if world > 1:
        dist.broadcast_object_list(box, src=0)
    s = pkg.Solver(local_rank)
    s.cascade_connect(rank, world, box[0])
    return s


def run_gpu(s, X, y, rank, world, mode, **params):
    """One cascade run from host buffers: partition upload + global statistics + rounds to convergence."""
    n = X.shape[0]
    start, end = partition(n, world, rank)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    s.cascade_set_partition_u8(X[start:end], y[start:end], start, n)
    out = s.cascade_train(mode, cache_rows=-1, **params)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    out["secs"] = time.perf_counter() - t0
    return out


def run_cpu_reference(pkg, X, digits, Xt, dgt, ranks: int, mode: str, timeout: float = 900.0):
    """The unmodified MPI program over the fork/pipe shim: `ranks` processes on the host cores."""
    if not os.path.exists(REF[mode]):
        return {"unavailable": "oracle/_ref cascade binaries not built"}
    with tempfile.TemporaryDirectory() as d:
        pkg.datagen.write_csv(os.path.join(d, "mnist3_train_data.csv"), X, digits)       # the names the reference hard-codes
        pkg.datagen.write_csv(os.path.join(d, "mnist3_test_data.csv"), Xt, dgt)
        t0 = time.perf_counter()
        r = subprocess.run([REF[mode]], cwd=d, env=dict(os.environ, MPI_SHIM_NP=str(ranks)), capture_output=True, text=True, timeout=timeout)
        wall = time.perf_counter() - t0
    if r.returncode != 0:
        return {"unavailable": f"exit code {r.returncode}: {r.stderr[-200:]}"}
    t = r.stdout

    def grab(pat, cast):
        m = re.search(pat, t)
        return cast(m.group(1)) if m else None
    return {"wall_secs": wall, "training_ms_reported": grab(r"training time = ([0-9.]+)", float), "b": grab(r"Final b = ([-0-9.]+)", float),
            "n_sv": grab(r"Final global SV count = (\d+)", int), "rounds": grab(r"Cascade finished in (\d+) rounds", int),
            "correct": grab(r"\((\d+)/\d+\)", int), "ranks": ranks}


def run(pkg, solver, X, y, rank, world, local_rank, args, digits=None):
    """The `cascade` block of bench.py's line (rank 0 returns it, the others None)."""
    import json
    n = X.shape[0]
    s = connect(pkg, local_rank, rank, world)
    gpath = os.path.join(ROOT, "tests", "golden", f"golden_n{n}.npz")
    g = np.load(gpath) if os.path.exists(gpath) else None
    Xt = yt = None
    if g is not None:
        Xt, dgt = pkg.datagen.make_dataset(int(g["m"]), split=1)
        yt = pkg.datagen.binary_labels(dgt)
    block = {"rows": n, "ranks": world, "what": "whole job from host buffers: partition upload + global min/max (ncclAllReduce) + rounds to "
             "convergence; SV tuples travel GPU to GPU (uint8 rows + 16-byte records) by ncclAllGather (star) / ncclSend+ncclRecv (tree) / ncclBroadcast"}
    modes = ["star"] + (["tree"] if world & (world - 1) == 0 else [])
    for mode in modes:
        run_gpu(s, X[: max(2000, world * 64)], y[: max(2000, world * 64)], rank, world, mode)           # warm-up (module load, NCCL channels)
        outs = [run_gpu(s, X, y, rank, world, mode) for _ in range(2)]         # the whole job twice; the faster run is reported, both are listed
        out = min(outs, key=lambda o: o["secs"])
        rec = {"secs": out["secs"], "secs_of_each_run": [o["secs"] for o in outs], "rounds": out["rounds"], "converged": bool(out["converged"]),
               "n_sv": out["n_sv"], "b": out["b"], "stages_on_rank0": out["stages"], "smo_iterations_on_rank0": out["smo_iterations"],
               "stage_secs_on_rank0": out["solve_ms"] / 1e3, "smo_kernel_secs_on_rank0": out["smo_ms"] / 1e3}
        if rank == 0 and g is not None:
            ids, _ = s.cascade_get_model()
            rec["sv_set_equals_undivided_reference"] = bool(np.array_equal(np.sort(ids), g["sv_ids"]))
            rec["b_rel_err_vs_undivided_reference"] = abs(out["b"] - float(g["b"])) / max(1.0, abs(float(g["b"])))
            rec["test_correct"] = int(s.predict_u8(Xt, yt, ge_zero=True)[0])
            rec["test_correct_undivided_reference"] = int(g["correct"])
        block[mode] = rec
    # the reference's MPI programs on the host cores, and the GPU cascade on the same bounded sample
    n_cpu = min(n, int(os.environ.get("SVMB200_CASCADE_CPU_ROWS", "8000")))
    ranks_cpu = world
    small = {}
    for mode in modes:
        small[mode] = run_gpu(s, X[:n_cpu], y[:n_cpu], rank, world, mode)
    if rank == 0 and not args.no_cpu:
        dg = digits if digits is not None else np.where(y > 0, 1, 0).astype(np.uint8)
        Xts, dgts = pkg.datagen.make_dataset(1000, split=1)
        cpu = {"rows": n_cpu, "ranks": ranks_cpu, "cores": os.cpu_count(),
               "what": "oracle/_ref/mpi_svm2 (star) and mpi_svm3 (tree): the reference's unmodified MPI programs over the fork/pipe MPI shim, one "
                       "process per rank on the host cores, reading the same rows as CSV; training_ms_reported is the program's own 'training time'"}
        for mode in modes:
            cpu[mode] = run_cpu_reference(pkg, X[:n_cpu], dg[:n_cpu], Xts, dgts, ranks_cpu, mode)
            cpu[mode]["gpu_secs_same_rows"] = small[mode]["secs"]
            cpu[mode]["gpu_n_sv_same_rows"] = small[mode]["n_sv"]
        block["cpu_reference"] = cpu
    s.close()
    return block if rank == 0 else None
