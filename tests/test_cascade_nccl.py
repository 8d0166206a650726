"""The device-resident Cascade SVM (svmb200_cascade_*: star = code/mpi_svm_main2.cpp:439-769, tree =
code/mpi_svm_main3.cpp:565-828) through the C ABI, pytest -m gpu.

With one rank (any GPU box) the cascade must arrive at the undivided problem's solution; with two ranks (needs two GPUs:
NCCL does not put two ranks on one device) the oracle is the reference itself -- the UNMODIFIED MPI programs over the
fork/pipe MPI shim (oracle/_ref/mpi_svm2, mpi_svm3) on the same rows: same SV count, rounds and accuracy, b within 1e-4."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from test_cascade import M, N, _reference

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _run(mode, world, tmp_path):
    out = tmp_path / f"nccl_{mode}_{world}.json"
    port = 29700 + (os.getpid() + hash((mode, world))) % 200
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "cascade_nccl_worker.py"), mode, str(N), str(M), str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert r.returncode == 0, r.stderr[-3000:]
    return json.load(open(out))


@pytest.mark.parametrize("mode", ["star", "tree"])
def test_single_rank_cascade_reaches_the_undivided_solution(mode, pkg, tmp_path):
    res = _run(mode, 1, tmp_path)[0]
    X, dg = pkg.datagen.make_dataset(N)
    Xt, dgt = pkg.datagen.make_dataset(M, split=1)
    s = pkg.Solver(0)
    try:
        s.set_train_u8(X, pkg.datagen.binary_labels(dg))
        whole = s.smo_train()
        assert res["converged"] and res["sv_ids"] == [int(i) for i in s.sv_ids()]
        assert abs(res["b"] - whole["b"]) <= 1e-4 * max(1.0, abs(whole["b"]))
        assert res["correct"] == s.predict_u8(Xt, pkg.datagen.binary_labels(dgt), ge_zero=True)[0]
    finally:
        s.close()


@pytest.mark.parametrize("mode", ["star", "tree"])
def test_two_gpu_cascade_matches_the_reference_programs(mode, pkg, tmp_path, tmp_path_factory):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    ref = _reference(mode, 2, pkg, tmp_path_factory)
    allres = _run(mode, 2, tmp_path)
    assert len(allres) == 2 and allres[0] == {**allres[1], "secs": allres[0]["secs"]}, "both ranks must hold the same final model"
    ours = allres[0]
    assert ours["converged"]
    assert len(ours["sv_ids"]) == ref["sv_count"]
    assert ours["rounds"] == ref["rounds"]
    assert ours["correct"] == ref["correct"]
    assert abs(ours["b"] - float(ref["b"])) <= 1e-4 * max(1.0, abs(float(ref["b"])))


@pytest.mark.parametrize("mode", ["star", "tree"])
def test_svm_cascade_program(mode, pkg, tmp_path, tmp_path_factory):
    """The C++ program that stands where mpi_svm2 / mpi_svm3 stand (one host thread per GPU): its report lines against the
    reference programs' on the same rows -- with every GPU of the box up to two ranks (one rank on a one-GPU box, where the
    comparison is with the reference run on one rank)."""
    import re
    pkgdir = os.path.dirname(pkg.__file__)
    subprocess.check_call(["make", "-s", "-C", pkgdir, "host"])
    world = min(2, _ngpu())
    ref = _reference(mode, world, pkg, tmp_path_factory)
    X, dg = pkg.datagen.make_dataset(N)
    Xt, dgt = pkg.datagen.make_dataset(M, split=1)
    pkg.datagen.write_idx(str(tmp_path / "tri"), str(tmp_path / "trl"), X, dg)
    pkg.datagen.write_idx(str(tmp_path / "tei"), str(tmp_path / "tel"), Xt, dgt)
    out = subprocess.run([os.path.join(pkgdir, "svm_cascade"), "--mode", mode, "--gpus", str(world), "--train-images", str(tmp_path / "tri"),
                          "--train-labels", str(tmp_path / "trl"), "--test-images", str(tmp_path / "tei"), "--test-labels", str(tmp_path / "tel")],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    t = out.stdout
    assert f"n = {N}" in t and "n_features = 784" in t
    assert int(re.search(r"Converged at round \d+, SV count = (\d+)", t).group(1)) == ref["sv_count"]
    assert int(re.search(r"Cascade finished in (\d+) rounds", t).group(1)) == ref["rounds"]
    assert int(re.search(r"\((\d+)/%d\)" % M, t).group(1)) == ref["correct"]
    b = float(re.search(r"Final b = ([-0-9.e+]+)", t).group(1))
    assert abs(b - float(ref["b"])) <= 1e-4 * max(1.0, abs(float(ref["b"])))
