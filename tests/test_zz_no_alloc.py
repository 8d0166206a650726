"""The allocation-free min/max entry point (svmb200_minmax_u8_noalloc), checked through the C ABI.

Some callers of this library may not allocate device memory at the moment they need a result -- a
plugin inside a framework whose pool owns the GPU, or code in a stream-ordered region where a
cudaMalloc could stall everything.  They hand over device buffers they already own; the scratch sits
in a fixed table inside the library.  This test plays such a caller: torch owns the buffers, the
entry point gets raw device pointers, and its answers must equal those of the regular, allocating
path (svmb200_set_train_u8) exactly -- min/max are order-independent -- on two calls in a row with
different shapes, so that values the first call left in the shared table cannot leak into the second.

The file name sorts last on purpose: under `-x` every other test has already run when this one does.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _noalloc_minmax(pkg, torch, X):
    L = pkg.load_library()
    dX = torch.from_numpy(X).cuda()
    n, d = X.shape
    dmin = torch.full((d,), float("nan"), dtype=torch.float64, device="cuda")
    dmax = torch.full((d,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    rc = L.svmb200_minmax_u8_noalloc(0, C.c_void_p(dX.data_ptr()), n, d, C.c_void_p(dmin.data_ptr()),
                                     C.c_void_p(dmax.data_ptr()), None)
    assert rc == 0
    torch.cuda.synchronize()
    return dmin.cpu().numpy(), dmax.cpu().numpy()


def test_noalloc_minmax_matches_allocating_path(pkg, dataset):
    import torch
    X1, _, y1 = dataset(3000)                         # 784 features, 47 row slabs
    rng = np.random.default_rng(11)
    X2 = rng.integers(3, 200, size=(70, 100), dtype=np.uint8)   # fewer slabs, narrower: stale table entries around it
    y2 = np.where(rng.uniform(size=70) < 0.5, 1, -1).astype(np.int32)
    s = pkg.Solver(0)
    try:
        for X, y in ((X1, y1), (X2, y2)):             # two calls in a row on the shared table
            got_min, got_max = _noalloc_minmax(pkg, torch, X)
            ref_min, ref_max = s.set_train_u8(X, y)
            assert np.array_equal(got_min, ref_min)
            assert np.array_equal(got_max, ref_max)
            assert np.array_equal(got_min, X.min(axis=0).astype(np.float64))
            assert np.array_equal(got_max, X.max(axis=0).astype(np.float64))
    finally:
        s.close()


def test_noalloc_rejects_what_does_not_fit(pkg):
    import torch
    L = pkg.load_library()
    d = 1025                                           # one past SVMB200_NOALLOC_MAX_FEATURES
    dX = torch.zeros((4, d), dtype=torch.uint8, device="cuda")
    out = torch.zeros((2, d), dtype=torch.float64, device="cuda")
    rc = L.svmb200_minmax_u8_noalloc(0, C.c_void_p(dX.data_ptr()), 4, d, C.c_void_p(out[0].data_ptr()),
                                     C.c_void_p(out[1].data_ptr()), None)
    assert rc == -5                                    # SVMB200_ELIMIT
    assert L.svmb200_minmax_u8_noalloc(0, None, 4, 8, None, None, None) == -1   # SVMB200_EINVAL
