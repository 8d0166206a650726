"""CPU tests of the drop-in boundary: the library loads and exports every symbol the header declares
(no compute without a GPU), and the product path fails loudly instead of falling back."""
import ctypes as C
import os
import subprocess

import pytest


def test_library_exports_every_declared_symbol(pkg):
    pkg.build_library()
    L = C.CDLL(pkg.LIB_PATH)
    names = pkg.declared_symbols()
    assert len(names) >= 15
    for name in names:
        assert hasattr(L, name), f"{name} declared in include/svmb200.h but not exported"
    nm = subprocess.run(["nm", "-D", "--defined-only", pkg.LIB_PATH], capture_output=True, text=True).stdout
    exported = {line.split()[-1] for line in nm.splitlines() if " T " in line}
    assert set(names) <= exported


def test_default_params_are_the_reference_literals(pkg):
    from_lib = pkg.binding.default_params()
    assert (from_lib.C, from_lib.gamma, from_lib.eps, from_lib.tau) == (10.0, 0.00125, 1e-12, 1e-5)
    assert (from_lib.max_iter, from_lib.sv_tol, from_lib.wss) == (100000, 1e-8, 1)


def test_no_cpu_fallback(pkg):
    """Without a GPU the product refuses to run; with one this test is vacuous."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.SvmB200Error):
        pkg.Solver(0)


def test_product_does_not_reference_the_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg_dir = os.path.join(root, "parallelizing-support-vector-machine-training-with-gpu-and-mpi_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_py" not in text and "liboracle" not in text and "libref_main3" not in text, f
