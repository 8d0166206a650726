"""GPU tests of context state handling and of the scalar step (pytest -m gpu), through the C ABI.

Covers: re-use of one context across training sets of different width / storage precision (the
support-vector block is re-sized for the new row pitch), save after load, the parameters a model
installed with set_model predicts with, and calculate_U_V + the pair update at the box corners
against the oracle (code/main3.cpp:145-159, :235-270)."""
import itertools

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _blobs(n, d, seed):
    rng = np.random.default_rng(seed)
    X = rng.integers(0, 256, size=(n, d)).astype(np.uint8)
    y = np.where(rng.random(n) < 0.5, 1, -1).astype(np.int32)
    return X, y


def test_context_reuse_from_narrow_to_wide_rows_and_to_exact_storage(pkg, O):
    """Every row becomes a support vector (random labels, large C), so the SV block of the second problem is far
    larger, in bytes, than what the first one allocated."""
    s = pkg.Solver(0)
    s.set_storage("f32")
    try:
        Xa, ya = _blobs(64, 16, 1)
        s.set_train_u8(Xa, ya)
        s.smo_train(C=1000.0, gamma=0.5)
        ca, _ = s.predict_u8(Xa, ya)
        assert ca == 64
        # wide rows, many support vectors
        Xb, yb = _blobs(1500, 784, 2)
        s.set_train_u8(Xb, yb)
        out = s.smo_train(C=1000.0, gamma=0.01)
        assert out["n_sv"] > 1200
        xs, mn, mx = O.scale(Xb)
        ref_c, ref_scores = O.predict(xs[:200], yb[:200], xs, yb, out["alpha"], out["b"], gamma=0.01, arith=O.ARITH_DEVICE)
        c, scores = s.predict_u8(Xb[:200], yb[:200])
        assert c == ref_c and np.array_equal(scores, ref_scores)
        # the same context switched to FP64 rows: the pitch doubles
        s.set_storage("f64")
        s.set_train_u8(Xb, yb)
        out = s.smo_train(C=1000.0, gamma=0.01)
        ref_c, ref_scores = O.predict(xs[:200], yb[:200], xs, yb, out["alpha"], out["b"], gamma=0.01, arith=O.ARITH_DEVICE_F64)
        c, scores = s.predict_u8(Xb[:200], yb[:200])
        assert c == ref_c and np.array_equal(scores, ref_scores)
    finally:
        s.close()


def test_save_after_load_round_trips(pkg, dataset, tmp_path):
    X, _, y = dataset(600)
    Xt, _, yt = dataset(300, split=1)
    s = pkg.Solver(0)
    t = pkg.Solver(0)
    u = pkg.Solver(0)
    try:
        s.set_train_u8(X, y)
        s.smo_train()
        c0, sc0 = s.predict_u8(Xt, yt)
        s.save_model(str(tmp_path / "a.model"))
        t.load_model(str(tmp_path / "a.model"))
        t.save_model(str(tmp_path / "b.model"))          # a loaded model has no training alphas: it saves its own copy
        assert (tmp_path / "a.model").read_bytes() == (tmp_path / "b.model").read_bytes()
        u.load_model(str(tmp_path / "b.model"))
        c2, sc2 = u.predict_u8(Xt, yt)
        assert c2 == c0 and np.array_equal(sc2, sc0)
    finally:
        s.close(); t.close(); u.close()


def test_set_model_predicts_with_the_trained_gamma(pkg, O, dataset):
    """cascade_main's sequence: train with a non-default gamma, upload another training set (the SV rows), install the
    model with set_model, predict: the kernel width must still be the trained one."""
    X, _, y = dataset(500)
    Xt, _, yt = dataset(200, split=1)
    gamma = 0.01
    s = pkg.Solver(0)
    try:
        xs, mn, mx = O.scale(X)
        xt, _, _ = O.scale(Xt, mn, mx)
        s.set_train_scaled(xs, y)
        out = s.smo_train(gamma=gamma)
        c_ref, sc_ref = s.predict_scaled(xt, yt, ge_zero=True)
        sv = s.sv_ids()
        s.set_train_scaled(xs[sv], y[sv])
        s.set_model(out["alpha"][sv], out["b"])
        c, sc = s.predict_scaled(xt, yt, ge_zero=True)
        assert c == c_ref and np.array_equal(sc, sc_ref)
    finally:
        s.close()


def test_pair_update_at_the_box_corners(solver, O):
    """calculate_U_V (main3.cpp:145-159), its feasibility test (:246), the eta guard (:253) and the clipped update
    (:259-266), evaluated on the device, against the oracle's restatement -- every sign combination with the alphas
    at 0, C and inside, including infeasible boxes and eta <= eps."""
    import ctypes as C
    Cc = 10.0
    alphas = [0.0, 1e-13, 2.5, Cc - 1e-13, Cc, Cc + 1.0, -1.0]
    cases = 0
    for yh, yl, ah, al in itertools.product((1, -1), (1, -1), alphas, alphas):
        for bh, bl, k12 in ((-1.0, 1.0, 0.3), (-0.2, 0.1, 0.999999), (0.5, 0.7, 1.0), (-3.0, 4.0, 0.0)):
            got, stop = solver.pair_update(yh, yl, ah, al, bh, bl, k12, Cc, 1e-12)
            s = yh * yl
            U, V = C.c_double(), C.c_double()
            feasible = O.lib().oracle_calculate_U_V(s, ah, al, Cc, C.byref(U), C.byref(V))
            assert got["U"] == U.value and got["V"] == V.value
            eta = 1.0 + 1.0 - 2.0 * k12
            assert got["eta"] == eta
            if not feasible:
                assert stop == 2 and got["alpha_high"] == ah and got["alpha_low"] == al
            elif eta <= 1e-12:
                assert stop == 3
            else:
                assert stop == 0
                nl = al + yl * (bh - bl) / eta
                nl = V.value if nl > V.value else nl
                nl = U.value if nl < U.value else nl
                nh = ah + s * (al - nl)
                assert got["alpha_low"] == nl and got["alpha_high"] == nh
                assert got["c_high"] == (nh - ah) * yh and got["c_low"] == (nl - al) * yl
            cases += 1
    assert cases == 4 * 49 * 4


def test_infeasible_box_stops_the_solver(pkg, O):
    """A warm start whose alphas lie outside [0, C] makes the first pair's box infeasible: the solver stops with
    SVMB200_STOP_INFEASIBLE (main3.cpp:246-251 prints and breaks) and still returns b, like the oracle."""
    rng = np.random.default_rng(5)
    X = rng.integers(0, 256, size=(40, 12)).astype(np.uint8)
    y = np.where(np.arange(40) % 2 == 0, 1, -1).astype(np.int32)
    a0 = np.zeros(40)
    a0[0] = 25.0          # y = +1, alpha > C: f_0 = 24 is the maximum over I_low; i_high is a y = +1 row with alpha = 0,
                          # so s = +1, U = max(0, 25 + 0 - C) = 15 > V = min(C, 25) = 10
    s = pkg.Solver(0)
    s.set_storage("f32")
    try:
        s.set_train_u8(X, y)
        out = s.smo_train(alpha0=a0, gamma=10.0)
        xs, _, _ = O.scale(X)
        ref = O.smo_train(xs, y, params=O.default_params(gamma=10.0), alpha0=a0, arith=O.ARITH_DEVICE)
        assert ref["stop_reason"] == 2, "the test's premise: the oracle hits the infeasible box"
        assert out["stop_reason"] == 2 and out["num_iter"] == ref["num_iter"]
        assert out["b"] == ref["b"] and np.array_equal(out["alpha"], ref["alpha"])
    finally:
        s.close()
