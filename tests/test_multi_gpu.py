"""Row-sharded SMO across GPUs (needs >= 2 GPUs: `gpurun --gpus 2`; skipped on a 1-GPU box)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_row_sharded_smo_is_bit_identical_to_one_gpu(world, tmp_path):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    out = tmp_path / "mgpu.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29400 + world), os.path.join(ROOT, "tests", "mgpu_worker.py"), "3000", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    allres = json.load(open(out))
    print(json.dumps(allres[0]))
    assert len(allres) == world
    for res in allres:
        for label in ("cache_off", "cache_on", "cache_small"):
            v = res[label]
            assert v["stop"] == 0 and v["iters"] == v["ref_iters"], (label, v)
            assert v["alpha_equal"] and v["b_equal"] and v["trace_equal"] and v["sv_equal"], (label, v)
        assert res["cache_on"]["cache_hits"] > 0 and res["cache_off"]["cache_hits"] == 0
    assert len({r["correct"] for r in allres}) == 1


@pytest.mark.gpu
@pytest.mark.parametrize("n", [40, 200])
def test_row_sharded_tiny_problems(n, tmp_path):
    """Grids of a few CTAs per GPU: fewer clusters than the 16 mailbox copies (one cluster at n = 40: a single copy; 6 at
    n = 200), same results as one GPU."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    out = tmp_path / "mgpu_tiny.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(29420 + n % 7), os.path.join(ROOT, "tests", "mgpu_worker.py"), str(n), str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    allres = json.load(open(out))
    assert len(allres) == 2
    for res in allres:
        for label in ("cache_off", "cache_on", "cache_small"):
            v = res[label]
            assert v["iters"] == v["ref_iters"], (label, v)
            assert v["alpha_equal"] and v["b_equal"] and v["trace_equal"] and v["sv_equal"], (label, v)


@pytest.mark.parametrize("world,n", [(2, 11), (3, 100)])
def test_row_sharding_host_logic_over_gloo(world, n, tmp_path):
    """CPU: the plumbing around the sharded solver (IPC-handle all_gather, row ranges, alpha all_gather, model install)."""
    out = tmp_path / "logic.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29300 + world), os.path.join(ROOT, "tests", "shard_logic_worker.py"), str(n), str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert r.returncode == 0, r.stderr[-3000:]
    assert json.load(open(out)) == [True] * world


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["star", "tree"])
def test_cascade_program_on_two_gpus(mode):
    """The program that stands where mpi_svm_main2/3 stand: two ranks, 4 000 synthetic rows; its report must name the
    SV count and accuracy of the undivided problem solved on one GPU."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    import re
    import numpy as np
    sys.path.insert(0, ROOT)
    import __graft_entry__ as entry
    pkg = entry.load_package()
    n = 4000
    X, dg = pkg.datagen.make_dataset(n); y = pkg.datagen.binary_labels(dg)
    Xt, dgt = pkg.datagen.make_dataset(10000, split=1); yt = pkg.datagen.binary_labels(dgt)
    s = pkg.Solver(0)
    try:
        s.set_train_u8(X, y)
        full = s.smo_train(cache_rows=-1)
        correct, _ = s.predict_u8(Xt, yt, ge_zero=True)
    finally:
        s.close()
    prog = os.path.join(ROOT, "parallelizing-support-vector-machine-training-with-gpu-and-mpi_b200", "cascade_main.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29411", prog, "--mode", mode, "--synthetic", str(n)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    out = r.stdout
    print(out[-800:])
    assert re.search(r"\[rank 0\] Converged at round \d+, SV count = (\d+)", out)
    assert int(re.search(r"Final global SV count = (\d+)", out).group(1)) == full["n_sv"]
    assert int(re.search(r"Test accuracy \(final model\) = [0-9.]+ \((\d+)/10000\)", out).group(1)) == correct
    b = float(re.search(r"Final b = ([-0-9.e]+)", out).group(1))
    assert abs(b - full["b"]) <= 1e-4 * max(1.0, abs(full["b"]))
