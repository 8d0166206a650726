"""Multi-GPU modes, one process per GPU, torch.distributed for the plumbing only.

* Row-sharded SMO (BASELINE.json configs[3]): `ShardedSolver`.  The per-iteration exchange happens
  inside the persistent kernels over NVLink peer memory (C ABI: svmb200_shard_*); torch.distributed
  only all-gathers the 64-byte IPC handles once and the alpha slices at the end.
* Cascade SVM, modified two-layer ("star", code/mpi_svm_main2.cpp:439-769) and classical binary tree
  (code/mpi_svm_main3.cpp:565-828): `cascade_star`, `cascade_tree`.  Each rank's sub-problem is solved
  by a `solve` callable -- the GPU solver in production; the CPU tests (gloo, world_size 2) inject the
  oracle to check the merge protocol itself.  The reference's rank-ordered MPI_Send/MPI_Recv gather
  (mpi_svm_main2.cpp:578-607) becomes one all_gather of counts plus one padded all_gather of the
  support-vector tuples; its Bcast of the merged model disappears because every rank solves the merged
  problem itself (same data, deterministic solver => same bits) -- see DESIGN.md 7.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")


# ------------------------------------------------------------------------------------------------
class ShardedSolver:
    """Row-sharded SMO over the ranks of the default process group (<= 8 GPUs of one box)."""

    def __init__(self, solver):
        self.s = solver
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        mine = torch.frombuffer(bytearray(solver.shard_setup(self.rank, self.world)), dtype=torch.uint8).to(_dev())
        allh = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(allh, mine)
        solver.shard_connect(self.world, b"".join(bytes(h.cpu().numpy().tobytes()) for h in allh))
        dist.barrier()

    def row_range(self, n):
        return n * self.rank // self.world, n * (self.rank + 1) // self.world

    def smo_train(self, **params):
        """All ranks call this in lockstep with the same parameters.  Returns the stats of this rank with
        the FULL alpha (all-gathered) and installs the model for sv_ids / predict."""
        n = self.s.n
        # the ranks meet here first: first-call module loads, a large cache allocation or a slow upload on one rank must not
        # eat into the ~3 s the kernels wait for a peer's record
        if dist.get_backend() == "nccl":
            torch.cuda.synchronize()
        dist.barrier()
        out = self.s.smo_train(**params)
        if out.get("stop_reason") == 5:
            raise RuntimeError(f"rank {self.rank}: a peer GPU's record never arrived (PEER_LOST); no model was produced")
        r0, r1 = self.row_range(n)
        chunk = (n + self.world - 1) // self.world + 1
        mine = torch.zeros(chunk, dtype=torch.float64)
        mine[: r1 - r0] = torch.from_numpy(out["alpha"][r0:r1])
        mine = mine.to(_dev())
        parts = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(parts, mine)
        alpha = np.zeros(n)
        for r, p in enumerate(parts):
            a0, a1 = n * r // self.world, n * (r + 1) // self.world
            alpha[a0:a1] = p.cpu().numpy()[: a1 - a0]
        out["alpha"] = alpha
        self.s.set_model(alpha, out["b"])
        out["n_sv"] = int((alpha > 1e-8).sum())
        return out


# ------------------------------------------------------------------------------------------------
# Cascade SVM.  A support-vector tuple is (X rows f64 [k,d], y int32 [k], alpha f64 [k], id int32 [k]),
# the 5-tuple (with the count) the reference ships between ranks (mpi_svm_main2.cpp:445-458,751-757).

def partition(n_total: int, world: int, rank: int):
    """Contiguous chunks of ceil(n/P) rows; global id = original row index (mpi_svm_main2.cpp:346-402)."""
    chunk = (n_total + world - 1) // world
    start = rank * chunk
    end = min((rank + 1) * chunk, n_total)
    return start, max(start, end)


def _extract_sv(X, y, alpha, ids, tol):
    idx = np.nonzero(alpha > tol)[0]                     # get_SV_indices, mpi_svm_main2.cpp:291-298
    return X[idx], y[idx], alpha[idx], ids[idx]


def _build_train_set(sv, X_part, y_part, id_part):
    """SV set first with its alpha (warm start), then the rows of the partition whose id is not among
    them, alpha = 0 (mpi_svm_main2.cpp:474-502 / mpi_svm_main3.cpp:629-655)."""
    Xs, ys, als, ids = sv
    seen = set(ids.tolist())
    keep = np.array([i for i, g in enumerate(id_part.tolist()) if not (g in seen or seen.add(g))], dtype=np.int64)
    X = np.concatenate([Xs, X_part[keep]]) if len(ids) else X_part[keep]
    y = np.concatenate([ys, y_part[keep]]) if len(ids) else y_part[keep]
    a = np.concatenate([als, np.zeros(len(keep))]) if len(ids) else np.zeros(len(keep))
    g = np.concatenate([ids, id_part[keep]]) if len(ids) else id_part[keep]
    return np.ascontiguousarray(X), np.ascontiguousarray(y, dtype=np.int32), a, np.ascontiguousarray(g, dtype=np.int32)


def _allgather_sv(sv, d):
    """All ranks receive every rank's tuple, in rank order: counts first, then one padded all_gather."""
    X, y, a, ids = sv
    dev = _dev()
    world = dist.get_world_size()
    cnt = torch.tensor([len(ids)], dtype=torch.int64, device=dev)
    cnts = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(cnts, cnt)
    cnts = [int(c.item()) for c in cnts]
    cap = max(max(cnts), 1)
    buf = torch.zeros((cap, d + 3), dtype=torch.float64)
    k = len(ids)
    if k:
        buf[:k, :d] = torch.from_numpy(np.ascontiguousarray(X, dtype=np.float64))
        buf[:k, d] = torch.from_numpy(y.astype(np.float64))
        buf[:k, d + 1] = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64))
        buf[:k, d + 2] = torch.from_numpy(ids.astype(np.float64))
    buf = buf.to(dev)
    bufs = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(bufs, buf)
    out = []
    for c, bb in zip(cnts, bufs):
        h = bb[:c].cpu().numpy()
        out.append((np.ascontiguousarray(h[:, :d]), h[:, d].astype(np.int32), np.ascontiguousarray(h[:, d + 1]), h[:, d + 2].astype(np.int32)))
    return out


def _sendrecv_sv(sv, d, dst=None, src=None):
    """Point-to-point tuple transfer for the tree merge (mpi_svm_main3.cpp:689-716)."""
    dev = _dev()
    if dst is not None:
        X, y, a, ids = sv
        k = len(ids)
        dist.send(torch.tensor([k], dtype=torch.int64, device=dev), dst)
        if k:
            buf = torch.zeros((k, d + 3), dtype=torch.float64)
            buf[:, :d] = torch.from_numpy(np.ascontiguousarray(X, dtype=np.float64))
            buf[:, d] = torch.from_numpy(y.astype(np.float64))
            buf[:, d + 1] = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64))
            buf[:, d + 2] = torch.from_numpy(ids.astype(np.float64))
            dist.send(buf.to(dev), dst)
        return None
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)
    dist.recv(cnt, src)
    k = int(cnt.item())
    if k == 0:
        return (np.zeros((0, d)), np.zeros(0, np.int32), np.zeros(0), np.zeros(0, np.int32))
    buf = torch.zeros((k, d + 3), dtype=torch.float64, device=dev)
    dist.recv(buf, src)
    h = buf.cpu().numpy()
    return (np.ascontiguousarray(h[:, :d]), h[:, d].astype(np.int32), np.ascontiguousarray(h[:, d + 1]), h[:, d + 2].astype(np.int32))


def _bcast_sv(sv, d, root=0):
    dev = _dev()
    rank = dist.get_rank()
    cnt = torch.tensor([len(sv[3]) if rank == root else 0], dtype=torch.int64, device=dev)
    dist.broadcast(cnt, root)
    k = int(cnt.item())
    if k == 0:
        return (np.zeros((0, d)), np.zeros(0, np.int32), np.zeros(0), np.zeros(0, np.int32))
    buf = torch.zeros((k, d + 3), dtype=torch.float64)
    if rank == root:
        X, y, a, ids = sv
        buf[:, :d] = torch.from_numpy(np.ascontiguousarray(X, dtype=np.float64))
        buf[:, d] = torch.from_numpy(y.astype(np.float64))
        buf[:, d + 1] = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64))
        buf[:, d + 2] = torch.from_numpy(ids.astype(np.float64))
    buf = buf.to(dev)
    dist.broadcast(buf, root)
    h = buf.cpu().numpy()
    return (np.ascontiguousarray(h[:, :d]), h[:, d].astype(np.int32), np.ascontiguousarray(h[:, d + 1]), h[:, d + 2].astype(np.int32))


def cascade_star(X_part, y_part, id_part, solve, tol_sv=1e-8, max_rounds=50, log=None):
    """Modified two-layer Cascade SVM (code/mpi_svm_main2.cpp:439-769).

    X_part: this rank's already-scaled rows (f64 [k, d]); id_part: their global ids.
    solve(X, y, alpha0) -> (alpha, b): the warm-started SMO_train(..., init=false) of the reference.
    Returns dict(sv=(X, y, alpha, id), b, rounds, converged) -- identical on every rank."""
    d = X_part.shape[1]
    empty = (np.zeros((0, d)), np.zeros(0, np.int32), np.zeros(0), np.zeros(0, np.int32))
    sv_global, b_global = empty, 0.0
    rnd, converged = 0, False
    while rnd < max_rounds and not converged:
        # layer 1: every rank trains on SV_global (alpha kept) + its partition (:474-507)
        X, y, a0, ids = _build_train_set(sv_global, X_part, y_part, id_part)
        if len(ids):
            alpha, _ = solve(X, y, a0)
            local_sv = _extract_sv(X, y, alpha, ids, tol_sv)                                     # :516-538
        else:
            local_sv = empty
        # gather: rank 0's SVs keep their alpha, the others join with alpha 0, first id wins (:552-607)
        allsv = _allgather_sv(local_sv, d)
        merged = allsv[0]
        for other in allsv[1:]:
            Xo, yo, _, io = other
            merged = _build_train_set(merged, Xo, yo, io)
        # layer 2: the merged problem (:615).  Every rank solves it: same inputs, deterministic solver.
        Xm, ym, am, im = merged
        if len(im):
            alpha_m, b_m = solve(Xm, ym, am)
            new_sv = _extract_sv(Xm, ym, alpha_m, im, tol_sv)                                    # :624-643
        else:
            new_sv, b_m = empty, 0.0
        same = len(new_sv[3]) == len(sv_global[3]) and set(new_sv[3].tolist()) == set(sv_global[3].tolist())   # :646-660
        sv_global, b_global = new_sv, b_m
        if log:
            log(f"round {rnd}: merged {len(im)} -> {len(new_sv[3])} SVs, converged={same}")
        converged = same
        rnd += 1
    return {"sv": sv_global, "b": b_global, "rounds": rnd, "converged": converged}


def cascade_tree(X_part, y_part, id_part, solve, tol_sv=1e-8, max_rounds=50, log=None):
    """Classical binary-tree Cascade SVM (code/mpi_svm_main3.cpp:565-828); world must be a power of two
    (:420-428).  The final model lives on rank 0 (as in the reference) and is broadcast on return."""
    rank, world = dist.get_rank(), dist.get_world_size()
    if world & (world - 1):
        raise ValueError("This cascade expects number of ranks to be a power of two.")     # mpi_svm_main3.cpp:425
    d = X_part.shape[1]
    empty = (np.zeros((0, d)), np.zeros(0, np.int32), np.zeros(0), np.zeros(0, np.int32))
    sv_mine = empty                      # X_sv, Y_sv, alpha_sv, ID_sv of the reference
    global_ids = np.zeros(0, np.int32)   # rank 0: the SV id set of the previous round (:560)
    b_local, rnd, converged = 0.0, 0, False
    while rnd < max_rounds:
        rnd += 1
        # rank 0 feeds its last result back to everyone (:574-613); round 1: empty
        recv = _bcast_sv(sv_mine if rank == 0 else empty, d, 0)
        cur = (X_part, y_part, np.zeros(len(id_part)), id_part)       # :616-619
        step = 1
        while step <= world:
            if rank % step == 0:                                       # :623-685
                X, y, a0, ids = _build_train_set(recv, cur[0], cur[1], cur[3])
                if len(ids):
                    alpha, b_local = solve(X, y, a0)
                    cur = _extract_sv(X, y, alpha, ids, tol_sv)
                else:
                    cur, b_local = empty, 0.0
            if step < world:                                           # :687-717
                if rank % (2 * step) == step:
                    _sendrecv_sv(cur, d, dst=rank - step)
                elif rank % (2 * step) == 0:
                    recv = _sendrecv_sv(None, d, src=rank + step)
            step *= 2
        sv_mine = cur
        flag = torch.zeros(1, dtype=torch.int64, device=_dev())
        if rank == 0:                                                  # :720-743
            same = len(global_ids) == len(cur[3]) and set(global_ids.tolist()) == set(cur[3].tolist())
            global_ids = cur[3].copy()
            flag[0] = 1 if same else 0
            if log:
                log(f"round {rnd}: {len(cur[3])} SVs, converged={same}")
        dist.broadcast(flag, 0)                                        # :823-825
        converged = bool(flag.item())
        if converged:
            break
    model = _bcast_sv(sv_mine if rank == 0 else empty, d, 0)
    bt = torch.tensor([b_local if rank == 0 else 0.0], dtype=torch.float64, device=_dev())
    dist.broadcast(bt, 0)
    return {"sv": model, "b": float(bt.item()), "rounds": rnd, "converged": converged}


def gpu_solve_fn(solver, **params):
    """solve(X, y, alpha0) over the C ABI: the reference's SMO_train(X, y, d, alpha, b, C, init=false)."""
    def solve(X, y, a0):
        solver.set_train_scaled(X, y)
        out = solver.smo_train(alpha0=a0, **params)
        return out["alpha"], out["b"]
    return solve
