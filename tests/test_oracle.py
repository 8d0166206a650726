"""CPU tests: the oracle restatement is pinned against the unmodified reference (oracle/_ref) and
against the committed golden fixtures; the data generator and file formats round-trip."""
import ctypes as C
import hashlib
import os

import numpy as np
import pytest

from conftest import golden


def _ref_or_skip(O):
    R = O.ref()
    if R is None:
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    return R


def test_kernel_matches_reference(O):
    R = _ref_or_skip(O)
    rng = np.random.default_rng(1)
    for d in (1, 4, 784):
        a, b = rng.uniform(0, 1, d), rng.uniform(0, 1, d)
        assert O.lib().oracle_kernel(a, b, d, 0.00125) == R.ref_kernel(a, b, d)
        assert O.lib().oracle_kernel(a, a, d, 0.00125) == 1.0  # exp(-0) (main3.cpp:239: K11)
    z = np.zeros(784)
    assert O.lib().oracle_kernel(z, z, 784, 0.00125) == 1.0


def test_scaling_matches_reference(O, dataset):
    R = _ref_or_skip(O)
    X, _, _ = dataset(300)
    x = X.astype(np.float64)
    mn, mx = np.zeros(784), np.zeros(784)
    R.ref_find_min_max(x, mn, mx, 300, 784)
    xs = x.copy()
    R.ref_scale_features(xs, mn, mx, 300, 784)
    xo, mno, mxo = O.scale(X)
    assert np.array_equal(mn, mno) and np.array_equal(mx, mxo)
    assert np.array_equal(xs, xo)
    # constant columns: range < 1e-12 -> 1 (main3.cpp:81-82) => scaled value 0
    const_cols = np.nonzero(mx - mn < 1e-12)[0]
    assert const_cols.size > 0 and np.all(xo[:, const_cols] == 0.0)


@pytest.mark.parametrize("n", [64, 300, 1000])
def test_smo_restatement_is_bit_identical_to_reference(O, dataset, n):
    R = _ref_or_skip(O)
    X, _, y = dataset(n)
    xs, _, _ = O.scale(X)
    a = np.zeros(n); b = C.c_double(); it = C.c_int()
    R.ref_smo_train(xs, y, n, 784, a, C.byref(b), 10.0, C.byref(it))
    o = O.smo_train(xs, y)
    assert np.array_equal(a, o["alpha"])
    assert b.value == o["b"]
    assert it.value == o["num_iter"]
    idx = np.zeros(n, dtype=np.int32)
    c = R.ref_sv_indices(a, n, 1e-8, idx)
    assert np.array_equal(idx[:c], O.sv_indices(o["alpha"]))


def test_bounded_reference_loop_matches_smo_train(O, dataset):
    """ref_smo_iterations (the timed CPU arm) walks the same trajectory as the reference's SMO_train."""
    R = _ref_or_skip(O)
    n = 400
    X, _, y = dataset(n)
    xs, _, _ = O.scale(X)
    full = O.smo_train(xs, y)
    a = np.zeros(n); rows = C.c_int64()
    done = R.ref_smo_iterations(xs, y, n, 784, 10.0, 10 ** 6, a.ctypes.data, C.byref(rows))
    assert done == full["num_iter"] - 1
    assert np.array_equal(a, full["alpha"])
    assert rows.value == full["rows_computed"]


@pytest.mark.parametrize("n", [256, 2000])
def test_oracle_reproduces_golden(O, pkg, dataset, n):
    g = golden(n)
    X, dg, y = dataset(n)
    assert hashlib.sha256(X.tobytes() + dg.tobytes()).hexdigest() == str(g["sha_train"]), "generator drifted"
    xs, mn, mx = O.scale(X)
    o = O.smo_train(xs, y, trace_cap=int(g["trace"].shape[0]))
    sv = O.sv_indices(o["alpha"])
    assert np.array_equal(sv, g["sv_ids"])
    assert np.array_equal(o["alpha"][sv], g["sv_alpha"])
    assert o["b"] == float(g["b"]) and o["num_iter"] == int(g["num_iter"])
    assert np.array_equal(o["trace"], g["trace"][:len(o["trace"])])
    m = int(g["m"])
    Xt, dgt, yt = dataset(m, split=1)
    assert hashlib.sha256(Xt.tobytes() + dgt.tobytes()).hexdigest() == str(g["sha_test"])
    xt, _, _ = O.scale(Xt, mn, mx)
    correct, scores = O.predict(xt, yt, xs, y, o["alpha"], o["b"])
    assert correct == int(g["correct"])
    assert np.array_equal(scores[:256], g["scores_head"][:min(256, m)])


def test_device_arithmetic_restated(O):
    """oracle_device_exp / oracle_device_kernel: close to libm / FP64 and exact on the diagonal."""
    L = O.lib()
    xs = np.concatenate([-np.logspace(-12, 2.8, 400), [0.0, -0.0, -745.0, -800.0]])
    for x in xs:
        e, r = L.oracle_device_exp(float(x)), float(np.exp(x))
        assert abs(e - r) <= 4e-16 * r + 5e-324, (x, e, r)
    assert L.oracle_device_exp(0.0) == 1.0
    rng = np.random.default_rng(3)
    for ch in (1, 2, 4):
        O.set_device_chunk(ch)
        for d in (3, 784):
            a = rng.uniform(0, 1, d).astype(np.float32)
            b = rng.uniform(0, 1, d).astype(np.float32)
            k = L.oracle_device_kernel(a, b, d, 0.00125)
            kr = L.oracle_kernel(a.astype(np.float64), b.astype(np.float64), d, 0.00125)
            assert abs(k - kr) <= 2e-7 * kr
            assert L.oracle_device_kernel(a, a, d, 0.00125) == 1.0
            assert L.oracle_device_kernel(a, b, d, 0.00125) == L.oracle_device_kernel(b, a, d, 0.00125)
    O.set_device_chunk(O.DEFAULT_CHUNK)


def test_selection_rules(O):
    """calc_i_high / calc_i_low: lowest index on ties, n when empty (main3.cpp:107-142)."""
    L = O.lib()
    n = 6
    y = np.array([1, 1, -1, -1, 1, -1], dtype=np.int32)
    alpha = np.zeros(n)
    f = -y.astype(np.float64)
    assert L.oracle_calc_i_high(alpha, y, f, n, 10.0, 1e-12) == 0     # y=+1, alpha<C : rows 0,1,4 tie at -1
    assert L.oracle_calc_i_low(alpha, y, f, n, 10.0, 1e-12) == 2      # y=-1, alpha<C : rows 2,3,5 tie at +1
    yp = np.ones(n, dtype=np.int32)
    assert L.oracle_calc_i_low(alpha, yp, -yp.astype(np.float64), n, 10.0, 1e-12) == n   # empty I_low
    U, V = C.c_double(), C.c_double()
    assert L.oracle_calculate_U_V(-1, 0.0, 0.0, 10.0, C.byref(U), C.byref(V)) == 1 and (U.value, V.value) == (0.0, 10.0)
    assert L.oracle_calculate_U_V(1, 10.0, 10.0, 10.0, C.byref(U), C.byref(V)) == 1 and (U.value, V.value) == (10.0, 10.0)
    assert L.oracle_calculate_U_V(1, 0.0, 0.0, 10.0, C.byref(U), C.byref(V)) == 1 and (U.value, V.value) == (0.0, 0.0)


def test_warm_start_restatement(O, dataset):
    """init=false (mpi_svm_main2.cpp:167-183): restarting from the solution stops at once with the same b."""
    n = 300
    X, _, y = dataset(n)
    xs, _, _ = O.scale(X)
    cold = O.smo_train(xs, y)
    warm = O.smo_train(xs, y, alpha0=cold["alpha"])
    assert warm["num_iter"] <= 3
    assert abs(warm["b"] - cold["b"]) < 1e-4
    assert np.array_equal(O.sv_indices(warm["alpha"]), O.sv_indices(cold["alpha"]))


def test_idx_and_csv_roundtrip(pkg, dataset, tmp_path):
    X, dg, _ = dataset(50)
    pkg.datagen.write_idx(str(tmp_path / "img"), str(tmp_path / "lab"), X, dg)
    X2, dg2 = pkg.datagen.read_idx(str(tmp_path / "img"), str(tmp_path / "lab"))
    assert np.array_equal(X, X2) and np.array_equal(dg, dg2)
    pkg.datagen.write_csv(str(tmp_path / "t.csv"), X, dg)
    rows = open(tmp_path / "t.csv").read().splitlines()
    assert len(rows) == 51 and rows[0].split(",")[-1] == "label"
    assert [int(v) for v in rows[1].split(",")] == X[0].tolist() + [int(dg[0])]


def test_dataset_prefix_property(pkg):
    a, da = pkg.datagen.make_dataset(5000)
    b, db = pkg.datagen.make_dataset(300)
    assert np.array_equal(a[:300], b) and np.array_equal(da[:300], db)
    nz = (a > 0).mean()
    assert 0.1 < nz < 0.35


@pytest.mark.parametrize("n", [256, 2000])
def test_u8_arithmetic_walks_the_reference_pair_sequence(O, dataset, n):
    """ORACLE_ARITH_DEVICE_U8 (exact integer sums over the features of the most frequent range, weighted FP64 sums over
    the rest) against the fixture of the unmodified reference: the kernel values agree to ~1e-16 and the solver selects the
    same pairs, iteration for iteration."""
    g = golden(n)
    X, _, y = dataset(n)
    R, M, T = O.u8_bind(X)
    assert R == 255 and M >= 32 and M + T <= 784
    xs, _, _ = O.scale(X)
    rng = np.random.default_rng(n)
    for _ in range(300):
        i, j = (int(v) for v in rng.integers(0, n, 2))
        ref = O.lib().oracle_kernel(xs[i], xs[j], 784, 0.00125)
        assert abs(O.u8_kernel(j, i) - ref) <= 4e-16 * ref
    assert O.u8_kernel(7, 7) == 1.0
    out = O.smo_train(xs, y, arith=O.ARITH_DEVICE_U8, trace_cap=4096)
    assert out["num_iter"] == int(g["num_iter"])
    assert np.array_equal(out["trace"], g["trace"][:len(out["trace"])])
    assert np.array_equal(O.sv_indices(out["alpha"]), g["sv_ids"])
    assert abs(out["b"] - float(g["b"])) <= 1e-12 * max(1.0, abs(float(g["b"])))


def test_second_order_selection_restatement(O, dataset):
    """wss = 2 has no counterpart in the reference (parity unpinned): it must reach the first-order rule's support-vector
    set and bias (to the stop tolerance) in fewer iterations, and never pick a pair the update would stop on."""
    X, _, y = dataset(1200)
    xs, _, _ = O.scale(X)
    one = O.smo_train(xs, y)
    two = O.smo_train(xs, y, params=O.default_params(wss=2), trace_cap=100000)
    assert two["stop_reason"] == 0 and two["num_iter"] < one["num_iter"]
    assert np.array_equal(O.sv_indices(two["alpha"]), O.sv_indices(one["alpha"]))
    assert abs(two["b"] - one["b"]) <= 1e-4 * max(1.0, abs(one["b"]))
    assert np.max(np.abs(two["alpha"] - one["alpha"])) <= 1e-3 * 10.0
    # the first pair: i_high as in the first-order rule, i_low by gain
    assert two["trace"][0][0] == one["i_high"] or two["trace"][0][0] == O.smo_train(xs, y, params=O.default_params(max_iter=1))["i_high"]


@pytest.mark.parametrize("n", [256, 2000])
def test_second_order_selection_agrees_with_libsvm(O, dataset, n):
    """The reference has no second-order rule, so its restatement is checked from outside as well: LIBSVM (through scikit-learn;
    its working-set selection is the rule of Fan, Chen, Lin 2005 that wss = 2 restates) on the same scaled rows, with its stop
    tolerance set to this solver's 2 tau, must end with the same support-vector set and bias."""
    svm = pytest.importorskip("sklearn.svm")
    X, _, y = dataset(n)
    xs, _, _ = O.scale(X)
    two = O.smo_train(xs, y, params=O.default_params(wss=2))
    clf = svm.SVC(C=10.0, kernel="rbf", gamma=0.00125, tol=2e-5, shrinking=False, cache_size=500).fit(xs, y)
    assert np.array_equal(np.sort(clf.support_), O.sv_indices(two["alpha"]))
    assert abs(-clf.intercept_[0] - two["b"]) <= 2e-4 * max(1.0, abs(two["b"]))      # decision = sum alpha y K - b here, + intercept there
    alpha_sk = np.zeros(len(y))
    alpha_sk[clf.support_] = np.abs(clf.dual_coef_[0])
    assert np.max(np.abs(alpha_sk - two["alpha"])) <= 5e-3 * 10.0                       # both stop at a gap of 2e-5: alphas agree to the tolerance
    assert abs(int(clf.n_iter_[0]) - (two["num_iter"] - 1)) <= 0.25 * two["num_iter"]   # the same rule: a similar number of steps
