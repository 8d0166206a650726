"""Second-order working-set selection (svmb200_params.wss = 2; Fan, Chen, Lin 2005), pytest -m gpu.

The reference has only the first-order rule (code/main3.cpp:107-142), so this rule's parity is UNPINNED by the reference:
the checker is the restatement in oracle/smo_oracle.c (oracle_params.wss = 2).  The GPU must walk that restatement's pair
sequence bit for bit in each storage's arithmetic, reach the same support-vector set as the first-order rule on the
fixtures, and do so in fewer iterations."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("storage,arith", [("f32", "ARITH_DEVICE"), ("f64", "ARITH_DEVICE_F64"), ("u8", "ARITH_DEVICE_U8")])
@pytest.mark.parametrize("n", [9, 256, 1500])
def test_wss2_bit_exact_vs_restatement(pkg, O, dataset, storage, arith, n):
    X, _, y = dataset(max(n, 16))
    X, y = X[:n], y[:n].copy()
    if n == 9:
        y[:] = [1, -1, -1, 1, -1, -1, -1, 1, -1]
    s = pkg.Solver(0)
    try:
        s.set_storage(storage)
        s.set_train_u8(X, y)
        xs, _, _ = O.scale(X)
        if storage == "u8":
            O.u8_bind(X)
        dev = O.smo_train(xs, y, params=O.default_params(wss=2), arith=getattr(O, arith), trace_cap=100000, want_f=True)
        first = O.smo_train(xs, y, arith=getattr(O, arith))
        outs = []
        for cache in (0, -1, 3):                                   # no cache: two X passes per iteration; whole / tiny cache
            out = s.smo_train(wss=2, cache_rows=cache, trace_cap=100000)
            assert out["num_iter"] == dev["num_iter"] and out["stop_reason"] == dev["stop_reason"], cache
            assert np.array_equal(s.get_trace(100000), dev["trace"]), cache
            assert np.array_equal(out["alpha"], dev["alpha"]) and out["b"] == dev["b"], cache
            assert np.array_equal(s.get_f(), dev["f"]), cache
            outs.append(out)
        assert outs[0]["cache_hits"] == 0 and outs[1]["x_passes"] < outs[0]["x_passes"]
        if n >= 256:
            assert dev["num_iter"] < first["num_iter"]
            assert np.array_equal(O.sv_indices(dev["alpha"]), O.sv_indices(first["alpha"]))
    finally:
        s.close()


@pytest.mark.parametrize("n", [2000, 10000])
def test_wss2_reaches_the_reference_solution(pkg, dataset, n):
    g = golden(n)
    X, _, y = dataset(n)
    Xt, _, yt = dataset(int(g["m"]), split=1)
    s = pkg.Solver(0)
    try:
        s.set_train_u8(X, y)
        out = s.smo_train(wss=2, cache_rows=-1)
        assert out["stop_reason"] == 0 and out["num_iter"] < int(g["num_iter"])
        assert np.array_equal(s.sv_ids(), g["sv_ids"])
        assert abs(out["b"] - float(g["b"])) <= 1e-4 * max(1.0, abs(float(g["b"])))
        assert s.predict_u8(Xt, yt)[0] == int(g["correct"])
    finally:
        s.close()
