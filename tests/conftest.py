import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def pkg():
    mod = entry.load_package()
    mod.build_library()          # no-op when libsvmb200.so is newer than its sources (nvcc cross-compiles without a GPU)
    return mod


@pytest.fixture(scope="session")
def O():
    o = entry.load_oracle()
    o.build()
    return o


@pytest.fixture(scope="session")
def solver(pkg):
    """One context on cuda:0 shared by the GPU tests (fails loudly without the built library).  fp32 rows: these tests
    check the fp32 arithmetic; the default storage of uint8 input (the bytes themselves) has tests/test_gpu_u8.py."""
    s = pkg.Solver(0)
    s.set_storage("f32")
    yield s
    s.close()


_DATA = {}


@pytest.fixture(scope="session")
def dataset(pkg):
    """dataset(n, split=0) -> (X uint8, digits, y +-1), cached."""
    def get(n, split=0, seed=587):
        key = (n, split, seed)
        if key not in _DATA:
            X, dg = pkg.datagen.make_dataset(n, seed=seed, split=split)
            _DATA[key] = (X, dg, pkg.datagen.binary_labels(dg))
        return _DATA[key]
    return get


def golden(n):
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden", f"golden_n{n}.npz")
    if not os.path.exists(path):
        pytest.skip(f"{path} not generated")
    return np.load(path, allow_pickle=False)
