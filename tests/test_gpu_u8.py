"""GPU parity tests of the uint8 storage of X (svmb200_set_storage "u8"; pytest -m gpu), through the C ABI.

Two levels, as in test_gpu_parity.py.  (1) Against ORACLE_ARITH_DEVICE_U8 -- the device arithmetic restated on
the CPU (exact integer sums over the features of the most frequent range, weighted FP64 sums over the rest) --
kernel rows, pair sequence, alpha, b and f must be BIT-EXACT.  (2) Against the golden fixtures generated from the
unmodified reference (code/main3.cpp): the SAME iteration count and (i_high, i_low) sequence, the same
support-vector ID set and test accuracy, b within 1e-11, alpha within 1e-7 relative (north_star asks for 1e-4)."""
import hashlib

import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


@pytest.fixture()
def u8_solver(pkg):
    s = pkg.Solver(0)
    s.set_storage("u8")
    yield s
    s.close()


def _oracle_u8(O, X, y, **kw):
    R, M, T = O.u8_bind(X)
    xs, mn, mx = O.scale(X)
    return O.smo_train(xs, y, arith=O.ARITH_DEVICE_U8, **kw), xs, (R, M, T), (mn, mx)


def test_layout_and_kernel_rows_bit_exact(u8_solver, O, dataset):
    n = 700
    X, _, y = dataset(n)
    X = X.copy()
    X[17] = X[3]          # duplicate rows -> K = 1 exactly
    X[40] = 0             # an all-zero row
    mn, mx = u8_solver.set_train_u8(X, y)
    R, M, T = O.u8_bind(X)
    xs, mno, mxo = O.scale(X)
    assert np.array_equal(mn, mno) and np.array_equal(mx, mxo)
    lay = u8_solver.u8_layout()
    assert (lay["main_range"], lay["main_features"], lay["tail_features"]) == (R, M, T)
    assert lay["row_bytes"] % 16 == 0 and lay["row_bytes"] >= 4 * ((M + 3) // 4 + (T + 3) // 4)
    assert M >= 32 and T > 0, "the fixture should exercise both the integer section and the weighted tail"
    # the scaled rows reconstructed from the bytes are the reference's scaled rows
    assert np.array_equal(u8_solver.get_x_scaled(), xs.astype(np.float32))
    for i in (0, 3, 40, 699):
        row = u8_solver.kernel_row(i)
        exp = np.array([O.u8_kernel(j, i) for j in range(n)])
        assert np.array_equal(row, exp), i
        ref = np.array([O.lib().oracle_kernel(xs[i], xs[j], 784, 0.00125) for j in range(n)])
        assert np.max(np.abs(row - ref) / ref) < 1e-15            # vs the reference's sequential FP64 sum + libm exp
    assert u8_solver.kernel_row(3)[17] == 1.0 and u8_solver.kernel_row(17)[3] == 1.0


@pytest.mark.parametrize("n", [1, 9, 256, 1000, 2000])
def test_training_bit_exact_vs_restated_arithmetic(u8_solver, O, dataset, n):
    X, _, y = dataset(max(n, 16))
    X, y = X[:n], y[:n].copy()
    if n == 9:
        y[:] = [1, -1, -1, 1, -1, -1, -1, 1, -1]
    u8_solver.set_train_u8(X, y)
    out = u8_solver.smo_train(trace_cap=100000)
    dev, xs, _, _ = _oracle_u8(O, X, y, trace_cap=100000, want_f=True)
    assert out["num_iter"] == dev["num_iter"] and out["stop_reason"] == dev["stop_reason"]
    assert np.array_equal(u8_solver.get_trace(100000), dev["trace"])
    assert np.array_equal(out["alpha"], dev["alpha"])
    assert out["b"] == dev["b"] and out["b_high"] == dev["b_high"] and out["b_low"] == dev["b_low"]
    assert (out["i_high"], out["i_low"]) == (dev["i_high"], dev["i_low"])
    assert np.array_equal(u8_solver.get_f(), dev["f"])
    assert np.array_equal(u8_solver.sv_ids(), O.sv_indices(dev["alpha"]))


@pytest.mark.parametrize("n", [256, 2000, 10000])
def test_walks_the_reference_pair_sequence(u8_solver, dataset, n):
    """The working-pair sequence IS the reference's (north_star), from the bytes themselves."""
    g = golden(n)
    X, dg, y = dataset(n)
    assert hashlib.sha256(X.tobytes() + dg.tobytes()).hexdigest() == str(g["sha_train"])
    u8_solver.set_train_u8(X, y)
    out = u8_solver.smo_train(trace_cap=4096, cache_rows=-1)
    assert out["num_iter"] == int(g["num_iter"])
    tr = u8_solver.get_trace(4096)
    assert np.array_equal(tr, g["trace"][:len(tr)])
    assert (out["i_high"], out["i_low"]) == (int(g["i_high"]), int(g["i_low"]))
    assert np.array_equal(u8_solver.sv_ids(), g["sv_ids"])
    assert abs(out["b"] - float(g["b"])) <= 1e-11 * max(1.0, abs(float(g["b"])))
    assert np.max(np.abs(out["alpha"][g["sv_ids"]] - g["sv_alpha"]) / g["sv_alpha"]) < 1e-7
    m = int(g["m"])
    Xt, _, yt = dataset(m, split=1)
    correct, scores = u8_solver.predict_u8(Xt, yt)
    assert correct == int(g["correct"])
    assert np.allclose(scores[:256], g["scores_head"][:min(256, m)], rtol=0, atol=1e-10)


def test_cache_warm_start_and_predict(u8_solver, O, dataset):
    n = 900
    X, _, y = dataset(n)
    Xt, _, yt = dataset(300, split=1)
    u8_solver.set_train_u8(X, y)
    off = u8_solver.smo_train(cache_rows=0, trace_cap=100000)
    tr_off = u8_solver.get_trace(100000)
    for cap in (-1, 16, 2):                                        # whole problem cached / FIFO eviction / thrashing
        on = u8_solver.smo_train(cache_rows=cap, trace_cap=100000)
        assert np.array_equal(on["alpha"], off["alpha"]) and on["b"] == off["b"] and on["num_iter"] == off["num_iter"], cap
        assert np.array_equal(u8_solver.get_trace(100000), tr_off)
        assert on["cache_hits"] + on["x_passes"] == on["num_iter"] - 1
    assert on["cache_hits"] >= 0 and u8_solver.smo_train(cache_rows=-1)["cache_hits"] > 0
    # the decision function works on scaled FP64 rows: bit-exact against the restated FP64-row arithmetic
    dev, xs, _, (mn, mx) = _oracle_u8(O, X, y)
    assert np.array_equal(dev["alpha"], off["alpha"]) and dev["b"] == off["b"]
    xt, _, _ = O.scale(Xt, mn, mx)
    c_dev, s_dev = O.predict(xt, yt, xs, y, off["alpha"], off["b"], arith=O.ARITH_DEVICE_U8)
    c_gpu, s_gpu = u8_solver.predict_u8(Xt, yt)
    assert c_gpu == c_dev and np.array_equal(s_gpu, s_dev)
    c_ref, s_ref = O.predict(xt, yt, xs, y, off["alpha"], off["b"])
    assert c_gpu == c_ref and np.max(np.abs(s_gpu - s_ref)) < 1e-10
    # warm start (init = false, mpi_svm_main2.cpp:167-183): f rebuilt from alpha, then the same loop
    part = O.smo_train(xs[:500], y[:500])
    a0 = np.zeros(n); a0[:500] = part["alpha"]
    O.u8_bind(X)
    devw = O.smo_train(xs, y, alpha0=a0, arith=O.ARITH_DEVICE_U8, trace_cap=100000, want_f=True)
    warm = u8_solver.smo_train(alpha0=a0, trace_cap=100000)
    assert warm["num_iter"] == devw["num_iter"]
    assert np.array_equal(u8_solver.get_trace(100000), devw["trace"])
    assert np.array_equal(warm["alpha"], devw["alpha"]) and warm["b"] == devw["b"]
    assert np.array_equal(u8_solver.get_f(), devw["f"])
    again = u8_solver.smo_train(alpha0=warm["alpha"])
    assert again["num_iter"] <= 3


def _ranged_bytes(rng, n, d, ranges):
    """uint8 rows whose feature k spans exactly [lo_k, lo_k + ranges[k]]."""
    lo = rng.integers(0, 256 - np.asarray(ranges), size=d)
    X = (lo + rng.integers(0, np.asarray(ranges) + 1, size=(n, d))).astype(np.uint8)
    for k in range(d):                                             # make sure both ends occur
        X[2 * k % n, k] = lo[k]
        X[(2 * k + 1) % n, k] = lo[k] + ranges[k]
    return X


def test_long_tails_in_short_tail_blocks(u8_solver, O):
    """832 features without a shared range: the tails of 32 rows do not fit beside nothing else in shared memory, so a warp's
    32-row block gets its tails in blocks of 16 rows whose sums are parked until the block ends; enough rows that every warp owns
    several such blocks.  The first iterations bit for bit against the restated arithmetic (kernel rows of both pivots enter f)."""
    d, n = 832, 24000
    rng = np.random.default_rng(832)
    ranges = np.asarray([1 + k % 254 for k in range(d)])
    X = _ranged_bytes(rng, n, d, ranges)
    w = rng.normal(size=d)
    score = X.astype(np.float64) @ w
    y = np.where(score > np.median(score), 1, -1).astype(np.int32)
    gamma = 0.5 / d
    u8_solver.set_train_u8(X, y)
    R, M, T = O.u8_bind(X)
    assert (M, T) == (0, d)
    out = u8_solver.smo_train(gamma=gamma, C=1.0, max_iter=12, trace_cap=64, cache_rows=0)
    xs, _, _ = O.scale(X)
    dev = O.smo_train(xs, y, params=O.default_params(gamma=gamma, C=1.0, max_iter=12), arith=O.ARITH_DEVICE_U8, trace_cap=64, want_f=True)
    assert out["num_iter"] == dev["num_iter"] and out["stop_reason"] == dev["stop_reason"] and dev["stop_reason"] in (4, "max_iter")
    assert np.array_equal(u8_solver.get_trace(64), dev["trace"])
    assert np.array_equal(out["alpha"], dev["alpha"]) and out["b"] == dev["b"]
    assert np.array_equal(u8_solver.get_f(), dev["f"])


@pytest.mark.parametrize("d,kind", [(1, "tail"), (4, "tail"), (30, "tail"), (100, "main"), (257, "mixed"), (600, "tail"),
                                    (770, "mixed"), (832, "main"), (832, "tail")])
def test_other_feature_counts_and_layouts(u8_solver, O, d, kind):
    """All-tail rows (no range shared by 32 features: up to 8 tail words per lane), all-main rows, mixtures with constant
    features; bit-exact against the restated arithmetic, same support vectors as the reference's FP64 arithmetic."""
    rng = np.random.default_rng(1000 + d)
    n = 300
    if kind == "tail":
        ranges = [1 + k % 254 for k in range(d)]                  # no range occurs 32 times: no integer section
    elif kind == "main":
        ranges = [200] * d
    else:
        ranges = [255 if k % 3 else (0 if k % 30 == 0 else 1 + k % 200) for k in range(d)]
    ranges = np.asarray(ranges)
    X = _ranged_bytes(rng, n, d, ranges)
    w = rng.normal(size=d)
    score = X.astype(np.float64) @ w
    y = np.where(score + rng.normal(0, 0.3 * score.std() + 1e-9, n) > np.median(score), 1, -1).astype(np.int32)
    if d == 1:
        y[::2] = 1; y[1::2] = -1
    gamma = 0.5 / d
    u8_solver.set_train_u8(X, y)
    R, M, T = O.u8_bind(X)
    lay = u8_solver.u8_layout()
    assert (lay["main_range"], lay["main_features"], lay["tail_features"]) == (R, M, T)
    if kind == "main":
        assert T == 0 and M == d
    if kind == "tail":
        assert M == 0 and T == d
    out = u8_solver.smo_train(gamma=gamma, C=1.0, trace_cap=100000)
    xs, mn, mx = O.scale(X)
    p = O.default_params(gamma=gamma, C=1.0)
    dev = O.smo_train(xs, y, params=p, arith=O.ARITH_DEVICE_U8, trace_cap=100000)
    assert out["num_iter"] == dev["num_iter"] and out["stop_reason"] == dev["stop_reason"]
    assert np.array_equal(u8_solver.get_trace(100000), dev["trace"])
    assert np.array_equal(out["alpha"], dev["alpha"]) and out["b"] == dev["b"]
    ref = O.smo_train(xs, y, params=p)
    assert np.array_equal(u8_solver.sv_ids(), O.sv_indices(ref["alpha"]))
    row = u8_solver.kernel_row(7, gamma=gamma)
    assert np.array_equal(row, np.array([O.u8_kernel(j, 7, gamma) for j in range(n)]))
    c_dev, s_dev = O.predict(xs[:77], y[:77], xs, y, out["alpha"], out["b"], gamma=gamma, arith=O.ARITH_DEVICE_U8)
    c_gpu, s_gpu = u8_solver.predict_u8(X[:77], y[:77])
    assert c_gpu == c_dev and np.array_equal(s_gpu, s_dev)


def test_one_vs_rest_labels_keep_the_cache(u8_solver, pkg, dataset):
    X, dg, _ = dataset(1200)
    u8_solver.set_train_u8(X, pkg.datagen.binary_labels(dg, 0))
    alone = []
    for digit in range(3):
        u8_solver.set_labels(pkg.datagen.binary_labels(dg, digit))
        alone.append(u8_solver.smo_train(cache_rows=-1))
    u8_solver.set_train_u8(X, pkg.datagen.binary_labels(dg, 0))
    for digit in range(3):
        u8_solver.set_labels(pkg.datagen.binary_labels(dg, digit))
        kept = u8_solver.smo_train(cache_rows=-1, keep_cache=1 if digit else 0)
        assert np.array_equal(kept["alpha"], alone[digit]["alpha"]) and kept["b"] == alone[digit]["b"]
        assert kept["x_passes"] <= alone[digit]["x_passes"]


def test_u8_storage_rejects_what_it_cannot_hold(pkg):
    s = pkg.Solver(0)
    try:
        s.set_storage("u8")
        y = np.array([1, -1, 1, -1], dtype=np.int32)
        with pytest.raises(pkg.SvmB200Error, match="uint8 storage"):
            s.set_train_f64(np.zeros((4, 3)), y)
        with pytest.raises(pkg.SvmB200Error, match="uint8 storage"):
            s.set_train_scaled(np.zeros((4, 3)), y)
        with pytest.raises(pkg.SvmB200Error, match="feature limit"):
            s.set_train_u8(np.zeros((4, 833), dtype=np.uint8), y)
        # all-constant features: every distance is 0, every kernel value 1 -> eta = 0 at the first pair
        s.set_train_u8(np.full((4, 5), 7, dtype=np.uint8), y)
        out = s.smo_train()
        assert out["stop_reason"] == 3
        assert s.u8_layout()["main_features"] == 0 and s.u8_layout()["tail_features"] == 0
        # back to fp32 storage on the same context
        s.set_storage("f32")
        s.set_train_f64(np.arange(12, dtype=np.float64).reshape(4, 3), y)
        s.smo_train()
    finally:
        s.close()


def test_full_size_60k_u8_is_the_reference_run(u8_solver, dataset):
    """configs[1] at BASELINE.json's size: the same iteration count, pair sequence, final pair, SV-ID set and accuracy as
    the unmodified reference; alpha and b far inside north_star's 1e-4."""
    n = 60000
    g = golden(n)
    X, dg, y = dataset(n)
    assert hashlib.sha256(X.tobytes() + dg.tobytes()).hexdigest() == str(g["sha_train"])
    u8_solver.set_train_u8(X, y)
    out = u8_solver.smo_train(cache_rows=0, trace_cap=4096)
    assert out["num_iter"] == int(g["num_iter"]) and out["stop_reason"] == 0
    assert np.array_equal(u8_solver.get_trace(4096), g["trace"][:4096])
    assert (out["i_high"], out["i_low"]) == (int(g["i_high"]), int(g["i_low"]))
    assert np.array_equal(u8_solver.sv_ids(), g["sv_ids"])
    assert abs(out["b"] - float(g["b"])) <= 1e-11
    assert np.max(np.abs(out["alpha"][g["sv_ids"]] - g["sv_alpha"]) / g["sv_alpha"]) < 1e-7
    Xt, _, yt = dataset(int(g["m"]), split=1)
    correct, _ = u8_solver.predict_u8(Xt, yt)
    assert correct == int(g["correct"])
    on = u8_solver.smo_train(cache_rows=-1)                        # the row cache changes no bit
    assert np.array_equal(on["alpha"], out["alpha"]) and on["b"] == out["b"] and on["num_iter"] == out["num_iter"]
