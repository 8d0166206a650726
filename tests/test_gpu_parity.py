"""GPU parity tests (pytest -m gpu): the CUDA path, called through the C ABI, against the oracle.

Two levels.  (1) Against ORACLE_ARITH_DEVICE -- the CUDA arithmetic restated on the CPU -- results
must be BIT-EXACT (kernel rows, pair sequence, alpha, b, f): any race or indexing bug shows up as a
difference, not as "numerical noise".  (2) Against ORACLE_ARITH_F64 / the golden fixtures from the
unmodified reference: identical support-vector ID set, identical test accuracy and iteration count
at the fixture sizes, b within 1e-4 relative, alpha within 1e-4 * C (tolerances per north_star;
fp32 X vs the reference's FP64 X)."""
import hashlib

import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu

B_RTOL = 1e-4          # relative to max(1, |b|): b = (b_high+b_low)/2 is only fixed to tau=1e-5 by the stop rule
ALPHA_ATOL_OVER_C = 1e-4


def _train_both(solver, O, X, y, chunk=4, **kw):
    solver.set_train_u8(X, y)
    out = solver.smo_train(accum_chunk=chunk, trace_cap=100000, **kw)
    xs, mn, mx = O.scale(X)
    O.set_device_chunk(chunk)
    dev = O.smo_train(xs, y, arith=O.ARITH_DEVICE, trace_cap=100000, want_f=True)
    O.set_device_chunk(O.DEFAULT_CHUNK)
    return out, dev, xs


def test_preprocessing_matches_oracle(solver, O, dataset):
    X, _, y = dataset(700)
    mn, mx = solver.set_train_u8(X, y)
    xs, mno, mxo = O.scale(X)
    assert np.array_equal(mn, mno) and np.array_equal(mx, mxo)
    assert np.array_equal(solver.get_x_scaled(), xs.astype(np.float32))
    # FP64 raw input (the reference's CSV path) gives the same rows
    mn2, mx2 = solver.set_train_f64(X.astype(np.float64), y)
    assert np.array_equal(mn2, mno) and np.array_equal(mx2, mxo)
    assert np.array_equal(solver.get_x_scaled(), xs.astype(np.float32))
    solver.set_train_scaled(xs, y)
    assert np.array_equal(solver.get_x_scaled(), xs.astype(np.float32))
    # the host-vector forms of find_min_max / scale_features: FP64 in, FP64 out, the reference's bits
    mn3, mx3 = solver.minmax_f64(X.astype(np.float64))
    assert np.array_equal(mn3, mno) and np.array_equal(mx3, mxo)
    assert np.array_equal(solver.scale_f64(X.astype(np.float64), mn3, mx3), xs)


@pytest.mark.parametrize("chunk", [1, 2, 4])
def test_kernel_row_bit_exact_vs_device_oracle(solver, O, dataset, chunk):
    X, _, y = dataset(600)
    X = X.copy()
    X[17] = X[3]          # duplicate rows -> K = 1 exactly
    X[40] = 0             # an all-zero row
    solver.set_train_u8(X, y)
    xs, _, _ = O.scale(X)
    x32 = xs.astype(np.float32)
    O.set_device_chunk(chunk)
    for i in (0, 3, 40, 599):
        row = solver.kernel_row(i, 0.00125, chunk)
        exp = np.array([O.lib().oracle_device_kernel(x32[i], x32[j], 784, 0.00125) for j in range(600)])
        assert np.array_equal(row, exp), f"row {i} chunk {chunk}"
        ref = np.array([O.lib().oracle_kernel(xs[i], xs[j], 784, 0.00125) for j in range(600)])
        assert np.max(np.abs(row - ref) / ref) < 2e-8
        assert row[i] == 1.0
    O.set_device_chunk(O.DEFAULT_CHUNK)
    assert solver.kernel_row(3)[17] == 1.0


def test_select_pair_rules(solver, O):
    rng = np.random.default_rng(5)
    L = O.lib()
    for n in (1, 7, 33, 1000, 5000):
        y = rng.choice([-1, 1], n).astype(np.int32)
        alpha = rng.choice([0.0, 10.0, 3.5, 1e-13], n)
        f = rng.choice([-1.0, 1.0, 0.25, np.inf, -np.inf], n)   # massive ties
        got = solver.select_pair(alpha, f, y)
        exp = (L.oracle_calc_i_high(alpha, y, f, n, 10.0, 1e-12), L.oracle_calc_i_low(alpha, y, f, n, 10.0, 1e-12))
        assert got == exp
    n = 100
    y = np.ones(n, dtype=np.int32)
    got = solver.select_pair(np.zeros(n), -y.astype(np.float64), y)
    assert got == (0, n)          # I_low empty -> n (main3.cpp:129)
    got = solver.select_pair(np.full(n, 10.0), -y.astype(np.float64), y)
    assert got == (n, 0)          # I_high empty


@pytest.mark.parametrize("n,chunk", [(1, 1), (9, 1), (256, 1), (256, 4), (1000, 2), (2000, 1)])
def test_training_bit_exact_vs_device_oracle(solver, O, dataset, n, chunk):
    X, _, y = dataset(max(n, 16))
    X, y = X[:n], y[:n].copy()
    if n == 9:
        y[:] = [1, -1, -1, 1, -1, -1, -1, 1, -1]
    out, dev, _ = _train_both(solver, O, X, y, chunk)
    assert out["num_iter"] == dev["num_iter"] and out["stop_reason"] == dev["stop_reason"]
    assert np.array_equal(solver.get_trace(100000), dev["trace"])
    assert np.array_equal(out["alpha"], dev["alpha"])
    assert out["b"] == dev["b"] and out["b_high"] == dev["b_high"] and out["b_low"] == dev["b_low"]
    assert (out["i_high"], out["i_low"]) == (dev["i_high"], dev["i_low"])
    assert np.array_equal(solver.get_f(), dev["f"])
    assert np.array_equal(solver.sv_ids(), O.sv_indices(dev["alpha"]))


@pytest.mark.parametrize("n", [256, 2000])
def test_training_matches_reference_golden(solver, O, dataset, n):
    g = golden(n)
    X, dg, y = dataset(n)
    assert hashlib.sha256(X.tobytes() + dg.tobytes()).hexdigest() == str(g["sha_train"])
    solver.set_train_u8(X, y)
    out = solver.smo_train(trace_cap=4096)
    assert np.array_equal(solver.sv_ids(), g["sv_ids"]), "support-vector ID set differs from the reference"
    assert abs(out["b"] - float(g["b"])) <= B_RTOL * max(1.0, abs(float(g["b"])))
    a_ref = np.zeros(n); a_ref[g["sv_ids"]] = g["sv_alpha"]
    assert np.max(np.abs(out["alpha"] - a_ref)) <= ALPHA_ATOL_OVER_C * 10.0
    # fp32 X against the reference's FP64 X: the pair sequence agrees for a long prefix and may then
    # part company at a near-tie (DESIGN.md "Numerics"); the optimum reached is the same
    assert abs(out["num_iter"] - int(g["num_iter"])) <= max(2, int(g["num_iter"]) // 50)
    tr = solver.get_trace(4096)
    k = min(len(tr), len(g["trace"]))
    diff = np.nonzero((tr[:k] != g["trace"][:k]).any(axis=1))[0]
    prefix = k if diff.size == 0 else int(diff[0])
    print(f"n={n}: iterations {out['num_iter']} vs reference {int(g['num_iter'])}; identical pair prefix {prefix}/{k}")
    assert prefix >= k // 2, "pair sequence leaves the reference's far too early"
    m = int(g["m"])
    Xt, _, yt = dataset(m, split=1)
    correct, scores = solver.predict_u8(Xt, yt)
    assert correct == int(g["correct"])
    assert np.allclose(scores[:256], g["scores_head"][:min(256, m)], rtol=0, atol=1e-4)


def test_cache_on_equals_cache_off(solver, O, dataset):
    X, _, y = dataset(1500)
    solver.set_train_u8(X, y)
    off = solver.smo_train(cache_rows=0, trace_cap=100000)
    tr_off = solver.get_trace(100000)
    f_off = solver.get_f()
    for cap in (-1, 64, 2):      # auto, small (FIFO eviction active), degenerate
        on = solver.smo_train(cache_rows=cap, trace_cap=100000)
        assert np.array_equal(on["alpha"], off["alpha"]) and on["b"] == off["b"] and on["num_iter"] == off["num_iter"]
        assert np.array_equal(solver.get_trace(100000), tr_off)
        assert np.array_equal(solver.get_f(), f_off)
        assert on["x_passes"] + on["cache_hits"] == on["num_iter"] - 1
        if cap == -1:
            assert on["cache_hits"] > 0.5 * on["num_iter"]
    assert off["cache_hits"] == 0 and off["x_passes"] == off["num_iter"] - 1
    # few rows per CTA and a cache far smaller than the working set: slots are recycled every few iterations, so a
    # K12 read from a recycled slot must not come from a stale L1 line of another SM
    X, _, y = dataset(300)
    solver.set_train_u8(X, y)
    off = solver.smo_train(cache_rows=0)
    for cap in (3, 8, 32):
        on = solver.smo_train(cache_rows=cap)
        assert np.array_equal(on["alpha"], off["alpha"]) and on["b"] == off["b"] and on["num_iter"] == off["num_iter"], cap


def test_same_cache_capacity_on_a_bigger_problem(solver, dataset):
    """An explicit cache capacity kept across uploads of growing size (the cache is sized in elements, not rows)."""
    for n in (200, 1200):
        X, _, y = dataset(n)
        solver.set_train_u8(X, y)
        off = solver.smo_train(cache_rows=0)
        on = solver.smo_train(cache_rows=48)
        assert np.array_equal(on["alpha"], off["alpha"]) and on["b"] == off["b"]


def test_one_vs_rest_with_a_kept_cache(solver, pkg, dataset):
    """Ten label vectors on one X: the kernel-row cache is label-independent, so keeping it changes no result."""
    X, dg, _ = dataset(1500)
    solver.set_train_u8(X, pkg.datagen.binary_labels(dg, 0))
    fresh, kept, passes = [], [], []
    for digit in range(4):
        solver.set_labels(pkg.datagen.binary_labels(dg, digit))
        fresh.append(solver.smo_train(cache_rows=0))
    solver.set_train_u8(X, pkg.datagen.binary_labels(dg, 0))
    for digit in range(4):
        solver.set_labels(pkg.datagen.binary_labels(dg, digit))
        o = solver.smo_train(cache_rows=-1, keep_cache=1)
        kept.append(o); passes.append(o["x_passes"])
    for a, b in zip(fresh, kept):
        assert np.array_equal(a["alpha"], b["alpha"]) and a["b"] == b["b"] and a["num_iter"] == b["num_iter"]
    alone = []
    for digit in range(4):
        solver.set_labels(pkg.datagen.binary_labels(dg, digit))
        alone.append(solver.smo_train(cache_rows=-1)["x_passes"])
    assert passes[0] == alone[0] and all(p <= q for p, q in zip(passes[1:], alone[1:])) and sum(passes[1:]) < sum(alone[1:])


def test_warm_start_matches_oracle(solver, O, dataset):
    """init=false: alpha in/out, f rebuilt from alpha (mpi_svm_main2.cpp:167-183)."""
    n = 800
    X, _, y = dataset(n)
    xs, _, _ = O.scale(X)
    # a feasible non-trivial start: the solution of the first 500 rows, zero elsewhere
    part = O.smo_train(xs[:500], y[:500])
    a0 = np.zeros(n); a0[:500] = part["alpha"]
    dev = O.smo_train(xs, y, alpha0=a0, arith=O.ARITH_DEVICE, trace_cap=100000, want_f=True)
    solver.set_train_u8(X, y)
    out = solver.smo_train(alpha0=a0, trace_cap=100000)
    assert out["num_iter"] == dev["num_iter"]
    assert np.array_equal(solver.get_trace(100000), dev["trace"])
    assert np.array_equal(out["alpha"], dev["alpha"]) and out["b"] == dev["b"]
    ref = O.smo_train(xs, y, alpha0=a0)
    assert np.array_equal(solver.sv_ids(), O.sv_indices(ref["alpha"]))
    assert abs(out["b"] - ref["b"]) <= B_RTOL * max(1.0, abs(ref["b"]))


def test_predict_matches_oracle(solver, O, dataset):
    n, m = 900, 400
    X, _, y = dataset(n)
    Xt, _, yt = dataset(m, split=1)
    solver.set_train_u8(X, y)
    out = solver.smo_train()
    xs, mn, mx = O.scale(X)
    xt, _, _ = O.scale(Xt, mn, mx)
    c_dev, s_dev = O.predict(xt, yt, xs, y, out["alpha"], out["b"], arith=O.ARITH_DEVICE)
    c_gpu, s_gpu = solver.predict_u8(Xt, yt)
    assert np.array_equal(s_gpu, s_dev) and c_gpu == c_dev            # bit-exact vs the restated arithmetic
    c_ref, s_ref = O.predict(xt, yt, xs, y, out["alpha"], out["b"])
    assert c_gpu == c_ref and np.max(np.abs(s_gpu - s_ref)) < 1e-5     # vs the reference's FP64 loop
    c_ge, _ = solver.predict_u8(Xt, yt, ge_zero=True)                   # MPI programs' ">= 0" rule
    assert c_ge == O.predict(xt, yt, xs, y, out["alpha"], out["b"], ge_zero=True, arith=O.ARITH_DEVICE)[0]
    c_f64, s_f64 = solver.predict_f64(Xt.astype(np.float64), yt)
    assert c_f64 == c_gpu and np.array_equal(s_f64, s_gpu)
    # a row's score does not depend on the batch it arrives in (ragged batches: 1, 33 and 383 rows; the kernel takes two
    # query rows per warp and picks its CTA size per launch)
    for k in (1, 33, 383):
        _, s_part = solver.predict_u8(Xt[:k], yt[:k])
        assert np.array_equal(s_part, s_gpu[:k]), k


def test_model_save_and_load(solver, pkg, dataset, tmp_path):
    """The model dump the reference left commented out (mpi_svm_main2.cpp:686-697), plus rows and statistics."""
    X, _, y = dataset(700)
    Xt, _, yt = dataset(300, split=1)
    solver.set_train_u8(X, y)
    out = solver.smo_train()
    c0, s0 = solver.predict_u8(Xt, yt)
    path = str(tmp_path / "model.svmb200")
    solver.save_model(path)
    fresh = pkg.Solver(0)
    try:
        fresh.load_model(path)
        assert np.array_equal(fresh.sv_ids(), solver.sv_ids())
        c1, s1 = fresh.predict_u8(Xt, yt)
        assert c1 == c0 and np.array_equal(s1, s0)
        with pytest.raises(pkg.SvmB200Error, match="before"):
            fresh.n = 700
            fresh.smo_train()
        open(path, "r+b").write(b"garbage!")
        with pytest.raises(pkg.SvmB200Error, match="not a svmb200 model"):
            fresh.load_model(path)
    finally:
        fresh.close()
    assert out["n_sv"] == len(solver.sv_ids())


@pytest.mark.parametrize("d", [1, 4, 30, 100, 257, 600, 770, 1024])
def test_other_feature_counts(solver, O, d):
    rng = np.random.default_rng(d)
    n = 300
    X = rng.uniform(-3, 5, (n, d))
    w = rng.normal(size=d)
    y = np.where(X @ w + rng.normal(0, 0.5, n) > np.median(X @ w), 1, -1).astype(np.int32)
    solver.set_train_f64(X, y)
    gamma = 0.5 / d
    out = solver.smo_train(gamma=gamma, C=1.0, trace_cap=100000)
    xs, _, _ = O.scale(X)
    p = O.default_params(gamma=gamma, C=1.0)
    dev = O.smo_train(xs, y, params=p, arith=O.ARITH_DEVICE, trace_cap=100000)
    assert out["num_iter"] == dev["num_iter"]
    assert np.array_equal(out["alpha"], dev["alpha"]) and out["b"] == dev["b"]
    ref = O.smo_train(xs, y, params=p)
    assert np.array_equal(solver.sv_ids(), O.sv_indices(ref["alpha"]))
    # the decision function at this shape (register groups partly or wholly beyond the row take other code paths)
    c_dev, s_dev = O.predict(xs[:77], y[:77], xs, y, out["alpha"], out["b"], gamma=gamma, arith=O.ARITH_DEVICE)
    c_gpu, s_gpu = solver.predict_f64(X[:77], y[:77])
    assert c_gpu == c_dev and np.array_equal(s_gpu, s_dev)


def test_abnormal_stops_are_stats_not_errors(solver, O, dataset, pkg):
    X, _, y = dataset(300)
    solver.set_train_u8(X, y)
    out = solver.smo_train(max_iter=10)
    xs, _, _ = O.scale(X)
    dev = O.smo_train(xs, y, params=O.default_params(max_iter=10), arith=O.ARITH_DEVICE)
    assert out["stop_reason"] == 4 == dev["stop_reason"] and out["num_iter"] == 11 == dev["num_iter"]
    assert np.array_equal(out["alpha"], dev["alpha"]) and out["b"] == dev["b"]
    yp = np.ones(300, dtype=np.int32)                        # one class only: I_low is empty at alpha = 0
    solver.set_train_u8(X, yp)
    out = solver.smo_train()
    assert out["stop_reason"] == 1 and out["num_iter"] == 1 and out["i_low"] == 300 and out["i_high"] == 0
    X2 = X.copy(); X2[1] = X2[0]                             # duplicate rows with opposite labels: eta = 0
    y2 = y.copy(); y2[:] = -1; y2[0] = 1; X2[2:] = X2[0]
    solver.set_train_u8(X2, y2)
    out = solver.smo_train()
    assert out["stop_reason"] == 3


def test_errors_are_reported(pkg):
    s = pkg.Solver(0)
    with pytest.raises(pkg.SvmB200Error, match="before"):
        s.smo_train()
    with pytest.raises(pkg.SvmB200Error):
        s.set_train_f64(np.zeros((4, 2000)), np.ones(4, dtype=np.int32))   # d > 1024
    X = np.zeros((8, 4)); y = np.array([1, -1] * 4, dtype=np.int32)
    s.set_train_f64(X, y)
    s.n = 8
    with pytest.raises(pkg.SvmB200Error, match="wss"):
        s.smo_train(wss=3)
    s.close()


# ---- the exact path: FP64 storage of X (svmb200_set_exact) -------------------------------------------------

@pytest.fixture()
def exact_solver(pkg):
    s = pkg.Solver(0)
    s.set_exact(True)
    yield s
    s.close()


def test_exact_path_bit_exact_vs_restated_arithmetic(exact_solver, O, dataset):
    n = 700
    X, _, y = dataset(n)
    X = X.copy(); X[11] = X[5]
    exact_solver.set_train_u8(X, y)
    xs, mn, mx = O.scale(X)
    for i in (0, 5, 699):
        row = exact_solver.kernel_row(i)
        exp = np.array([O.lib().oracle_device_kernel_f64(xs[i], xs[j], 784, 0.00125) for j in range(n)])
        assert np.array_equal(row, exp)
        ref = np.array([O.lib().oracle_kernel(xs[i], xs[j], 784, 0.00125) for j in range(n)])
        assert np.max(np.abs(row - ref) / ref) < 1e-15            # vs the reference's sequential FP64 sum + libm exp
    assert exact_solver.kernel_row(5)[11] == 1.0
    out = exact_solver.smo_train(trace_cap=100000)
    dev = O.smo_train(xs, y, arith=O.ARITH_DEVICE_F64, trace_cap=100000, want_f=True)
    assert out["num_iter"] == dev["num_iter"]
    assert np.array_equal(exact_solver.get_trace(100000), dev["trace"])
    assert np.array_equal(out["alpha"], dev["alpha"]) and out["b"] == dev["b"]
    assert np.array_equal(exact_solver.get_f(), dev["f"])
    Xt, _, yt = dataset(200, split=1)
    xt, _, _ = O.scale(Xt, mn, mx)
    c_dev, s_dev = O.predict(xt, yt, xs, y, out["alpha"], out["b"], arith=O.ARITH_DEVICE_F64)
    c_gpu, s_gpu = exact_solver.predict_u8(Xt, yt)
    assert c_gpu == c_dev and np.array_equal(s_gpu, s_dev)
    for cap in (-1, 16):                                           # row cache on the exact path
        on = exact_solver.smo_train(cache_rows=cap)
        assert np.array_equal(on["alpha"], out["alpha"]) and on["b"] == out["b"]
    warm = exact_solver.smo_train(alpha0=out["alpha"])
    assert warm["num_iter"] <= 3


@pytest.mark.parametrize("n", [256, 2000, 10000])
def test_exact_path_walks_the_reference_pair_sequence(exact_solver, dataset, n):
    """FP64 rows: the working-pair sequence IS the reference's (north_star), not just a prefix of it."""
    g = golden(n)
    X, _, y = dataset(n)
    exact_solver.set_train_u8(X, y)
    out = exact_solver.smo_train(trace_cap=4096, cache_rows=-1)
    assert out["num_iter"] == int(g["num_iter"])
    tr = exact_solver.get_trace(4096)
    assert np.array_equal(tr, g["trace"][:len(tr)])
    assert (out["i_high"], out["i_low"]) == (int(g["i_high"]), int(g["i_low"]))
    assert np.array_equal(exact_solver.sv_ids(), g["sv_ids"])
    assert abs(out["b"] - float(g["b"])) <= 1e-12 * max(1.0, abs(float(g["b"])))
    assert np.max(np.abs(out["alpha"][g["sv_ids"]] - g["sv_alpha"]) / g["sv_alpha"]) < 1e-9
    m = int(g["m"])
    Xt, _, yt = dataset(m, split=1)
    correct, scores = exact_solver.predict_u8(Xt, yt)
    assert correct == int(g["correct"])
    assert np.allclose(scores[:256], g["scores_head"][:min(256, m)], rtol=0, atol=1e-11)


# ---- BASELINE.json's full size (60 000 x 784) ------------------------------------------------------------------

def test_full_size_60k_against_reference_fixture_and_invariants(solver, dataset):
    """configs[1]: the 60k problem.  Fixture from the reference's algorithm (tests/golden/golden_n60000.npz) plus
    size-independent properties of an SMO solution checked from the returned state alone."""
    n = 60000
    g = golden(n)
    X, dg, y = dataset(n)
    assert hashlib.sha256(X.tobytes() + dg.tobytes()).hexdigest() == str(g["sha_train"])
    solver.set_train_u8(X, y)
    out = solver.smo_train(cache_rows=-1)
    alpha, f = out["alpha"], solver.get_f()
    # the reference's result
    assert np.array_equal(solver.sv_ids(), g["sv_ids"])
    assert abs(out["b"] - float(g["b"])) <= B_RTOL * max(1.0, abs(float(g["b"])))
    a_ref = np.zeros(n); a_ref[g["sv_ids"]] = g["sv_alpha"]
    # fp32 rows at full size: after ~19 000 iterations the pair sequence has long left the reference's, and the stop rule
    # (b_low <= b_high + 2e-5) under-determines alpha: 1e-4 holds in the L2 sense, single alphas move by up to ~5e-4 C.
    # The exact path below meets the element-wise bound.
    assert np.linalg.norm(alpha - a_ref) <= 1e-4 * np.linalg.norm(a_ref)
    assert np.max(np.abs(alpha - a_ref)) <= 1e-3 * 10.0
    assert abs(out["num_iter"] - int(g["num_iter"])) <= int(g["num_iter"]) // 50
    Xt, _, yt = dataset(int(g["m"]), split=1)
    correct, _ = solver.predict_u8(Xt, yt)
    assert correct == int(g["correct"])
    # invariants: box, equality constraint, stop rule (main3.cpp:213) on the final f, pair optimality
    C_, eps, tau = 10.0, 1e-12, 1e-5
    assert alpha.min() >= -1e-12 and alpha.max() <= C_ + 1e-12     # only alpha_low is clipped (main3.cpp:261-266): alpha_high rounds
    assert abs(float(np.dot(alpha, y))) <= 1e-9 * float(np.abs(alpha).sum())
    in_high = ((y == 1) & (alpha < C_ - eps)) | ((y == -1) & (alpha > eps))
    in_low = ((y == 1) & (alpha > eps)) | ((y == -1) & (alpha < C_ - eps))
    b_high, b_low = f[in_high].min(), f[in_low].max()
    assert b_low <= b_high + 2 * tau and out["stop_reason"] == 0
    assert (out["b_high"], out["b_low"]) == (b_high, b_low) and out["b"] == (b_high + b_low) / 2
    assert (out["i_high"], out["i_low"]) == (int(np.nonzero(in_high & (f == b_high))[0][0]), int(np.nonzero(in_low & (f == b_low))[0][0]))
    # f was integrated over ~19 000 updates: it must still be sum_j alpha_j y_j K(i,j) - y_i
    sv = np.nonzero(alpha > 0)[0]
    for i in (0, 12345, 59999, int(sv[0]), int(sv[-1])):
        k = solver.kernel_row(i)
        assert abs(float(np.dot(alpha[sv] * y[sv], k[sv])) - y[i] - f[i]) < 1e-8
    # memory-side choices change no bit: the L2-resident share of X off / on, the row cache off / on
    # (only at this size do the rows exceed L2, so only here is the residency policy in force)
    runs = [solver.smo_train(max_iter=1500, cache_rows=c, l2_persist=p, trace_cap=1500) for c, p in ((0, 0), (0, 1), (-1, 1))]
    for r in runs[1:]:
        assert np.array_equal(r["alpha"], runs[0]["alpha"]) and r["b"] == runs[0]["b"] and r["num_iter"] == runs[0]["num_iter"]


def test_full_size_60k_exact_path_is_the_reference_run(exact_solver, dataset):
    n = 60000
    g = golden(n)
    X, _, y = dataset(n)
    exact_solver.set_train_u8(X, y)
    out = exact_solver.smo_train(cache_rows=-1, trace_cap=4096)
    assert out["num_iter"] == int(g["num_iter"])
    assert np.array_equal(exact_solver.get_trace(4096), g["trace"][:4096])
    assert (out["i_high"], out["i_low"]) == (int(g["i_high"]), int(g["i_low"]))
    assert np.array_equal(exact_solver.sv_ids(), g["sv_ids"])
    assert abs(out["b"] - float(g["b"])) <= 1e-11
    assert np.max(np.abs(out["alpha"][g["sv_ids"]] - g["sv_alpha"]) / g["sv_alpha"]) < 1e-7   # north_star asks 1e-4


def test_size_limits_and_degenerate_inputs(pkg):
    s = pkg.Solver(0)
    try:
        y2 = np.array([1, -1], dtype=np.int32)
        with pytest.raises(pkg.SvmB200Error, match="positive"):
            s.set_train_f64(np.zeros((0, 4)), np.zeros(0, dtype=np.int32))                 # empty training set
        s.set_train_f64(np.zeros((2, 1024)), y2)                                             # widest fp32 rows
        with pytest.raises(pkg.SvmB200Error, match="1024"):
            s.set_train_f64(np.zeros((2, 1025)), y2)
        s.set_exact(True)
        s.set_train_f64(np.zeros((2, 832)), y2)                                              # widest exact rows
        with pytest.raises(pkg.SvmB200Error, match="832"):
            s.set_train_f64(np.zeros((2, 833)), y2)
        s.set_exact(False)
        X = np.array([[0.0, 1.0], [1.0, 0.0]])                                               # the smallest separable problem
        s.set_train_f64(X, y2)
        out = s.smo_train()
        assert out["stop_reason"] == 0 and out["n_sv"] == 2 and abs(out["alpha"][0] - out["alpha"][1]) < 1e-12
        c, sc = s.predict_f64(X, y2)
        assert c == 2 and sc[0] > 0 > sc[1]
    finally:
        s.close()


# ---- the decision function on the tensor cores (tcgen05, TF32x3) --------------------------------------------------

def test_tensor_core_dot_tile(solver, dataset):
    X, _, y = dataset(300)
    solver.set_train_u8(X, y)
    xs = solver.get_x_scaled()[:128].astype(np.float64)
    ref = xs @ xs.T
    got = solver.tc_dots().astype(np.float64)
    err = np.abs(got - ref)
    # three tf32 products per element, ~2 400 fp32 accumulations per dot product (the tensor core truncates them); |dot| <= ~150
    assert err.max() <= 4e-3 and np.median(err) <= 5e-4, (err.max(), np.median(err))
    assert np.abs(got - got.T).max() <= 4e-3


def test_tensor_core_predict_equals_exact_labels(solver, O, dataset):
    n, m = 3000, 1500
    X, _, y = dataset(n)
    Xt, _, yt = dataset(m, split=1)
    solver.set_train_u8(X, y)
    solver.smo_train(cache_rows=-1)
    c_exact, s_exact = solver.predict_u8(Xt, yt)
    try:
        solver.set_tensor_cores(True, 0.05)
        c_tc, s_tc = solver.predict_u8(Xt, yt)
        rechecked = solver.tc_rechecked()
    finally:
        solver.set_tensor_cores(False)
    assert c_tc == c_exact
    assert np.array_equal(s_tc > 0, s_exact > 0)
    assert np.max(np.abs(s_tc - s_exact)) < 2e-2                  # the tensor-core scores themselves: 2.5x under the threshold
    unc = np.abs(s_tc) < 0.05
    assert np.array_equal(s_tc[unc], s_exact[unc])                # rows under the threshold were re-evaluated exactly
    assert 0 <= rechecked <= m // 4
    # a threshold above every |score|: all rows go through the re-evaluation (one CTA per row instead of several) and
    # every score is the exact kernel's, bit for bit; then a batch of ONE row (a single pair of tiles)
    try:
        solver.set_tensor_cores(True, 1e3)
        c_all, s_all = solver.predict_u8(Xt, yt)
        assert solver.tc_rechecked() == m
        c_one, s_one = solver.predict_u8(Xt[:1], yt[:1])
    finally:
        solver.set_tensor_cores(False)
    assert c_all == c_exact and np.array_equal(s_all, s_exact)
    assert np.array_equal(s_one, s_exact[:1])
    print(f"tensor-core predict: max |score error| {np.max(np.abs(s_tc - s_exact)):.2e}, {rechecked} of {m} rows re-evaluated")
