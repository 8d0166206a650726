"""The C++ host layer (host/svmb200_host.*) and the svm_train program: the reference's own call sequence
(code/main3.cpp:306-417) compiled against svmb200 and compared with the oracle."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "parallelizing-support-vector-machine-training-with-gpu-and-mpi_b200")


def _build_host():
    subprocess.check_call(["make", "-s", "-C", PKG, "host"])


def test_host_layer_builds_and_exports_reference_names():
    _build_host()
    nm = subprocess.run(["nm", "-DC", "--defined-only", os.path.join(PKG, "libsvmb200_host.so")], capture_output=True, text=True).stdout
    for name in ("svmb200::SMO_train(", "svmb200::find_min_max(", "svmb200::scale_features(", "svmb200::get_SV_indices(",
                 "svmb200::read_CSV(", "svmb200::read_idx(", "svmb200::predict("):
        assert name in nm, name
    out = subprocess.run([os.path.join(PKG, "svm_train"), "--help"], capture_output=True, text=True)
    assert out.returncode == 0 and "usage: svm_train" in out.stderr


@pytest.mark.gpu
def test_reference_main_flow_over_host_layer(pkg, O, dataset, tmp_path):
    _build_host()
    n, m = 700, 300
    X, dg, y = dataset(n)
    Xt, dgt, yt = dataset(m, split=1)
    pkg.datagen.write_csv(str(tmp_path / "syn_train_data.csv"), X, dg)
    pkg.datagen.write_csv(str(tmp_path / "syn_test_data.csv"), Xt, dgt)
    exe = str(tmp_path / "main3_flow")
    subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O2", "-I", os.path.join(PKG, "host"), "-o", exe,
                           os.path.join(ROOT, "tests", "host_main3_flow.cpp"), "-L", PKG, "-lsvmb200_host", "-lsvmb200",
                           f"-Wl,-rpath,{PKG}"])
    out = subprocess.run([exe, "syn"], cwd=tmp_path, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    txt = out.stdout
    xs, mn, mx = O.scale(X)
    xt, _, _ = O.scale(Xt, mn, mx)
    ref = O.smo_train(xs, y)
    sv_ref = O.sv_indices(ref["alpha"])
    c_ref, _ = O.predict(xt, yt, xs, y, ref["alpha"], ref["b"])
    assert f"n = {n}" in txt and "n_features = 784" in txt
    b = float(re.search(r"^b = ([-0-9.e+]+)", txt, re.M).group(1))
    assert abs(b - ref["b"]) <= 1e-4 * max(1.0, abs(ref["b"]))
    assert int(re.search(r"Final SV count = (\d+)", txt).group(1)) == len(sv_ref)
    ids = np.array(re.search(r"SV ids:(.*)", txt).group(1).split(), dtype=np.int64)
    assert np.array_equal(ids, sv_ref)
    assert int(re.search(r"Test accuracy = [0-9.]+ \((\d+)/", txt).group(1)) == c_ref
    its = [int(v) for v in re.findall(r"number of iterations: (\d+)", txt)]
    assert abs(its[0] - ref["num_iter"]) <= max(2, ref["num_iter"] // 50) and its[1] <= 3   # the warm restart stops at once
    b2 = float(re.search(r"warm restart b = ([-0-9.e+]+)", txt).group(1))
    assert abs(b2 - ref["b"]) <= 1e-4 * max(1.0, abs(ref["b"]))


@pytest.mark.gpu
def test_svm_train_program_on_idx_files(pkg, O, dataset, tmp_path):
    _build_host()
    n, m = 900, 300
    X, dg, y = dataset(n)
    Xt, dgt, yt = dataset(m, split=1)
    pkg.datagen.write_idx(str(tmp_path / "tri"), str(tmp_path / "trl"), X, dg)
    pkg.datagen.write_idx(str(tmp_path / "tei"), str(tmp_path / "tel"), Xt, dgt)
    out = subprocess.run([os.path.join(PKG, "svm_train"), "--train-images", str(tmp_path / "tri"), "--train-labels", str(tmp_path / "trl"),
                          "--test-images", str(tmp_path / "tei"), "--test-labels", str(tmp_path / "tel"), "--limit", "800"],
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    xs, mn, mx = O.scale(X[:800])
    xt, _, _ = O.scale(Xt, mn, mx)
    ref = O.smo_train(xs, y[:800])
    c_ref, _ = O.predict(xt, yt, xs, y[:800], ref["alpha"], ref["b"])
    assert "n = 800" in out.stdout
    assert int(re.search(r"Final SV count = (\d+)", out.stdout).group(1)) == len(O.sv_indices(ref["alpha"]))
    assert int(re.search(r"Test accuracy = [0-9.]+ \((\d+)/300\)", out.stdout).group(1)) == c_ref
    b = float(re.search(r"^b = ([-0-9.e+]+)", out.stdout, re.M).group(1))
    assert abs(b - ref["b"]) <= 1e-4 * max(1.0, abs(ref["b"]))
