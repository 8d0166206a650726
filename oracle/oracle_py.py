"""ctypes loader for the CPU checker (oracle/liboracle.so, oracle/_ref/libref_main3.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; never by the product path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ARITH_F64, ARITH_DEVICE, ARITH_DEVICE_F64, ARITH_DEVICE_U8 = 0, 1, 2, 3
DEFAULT_CHUNK = 4   # the library's default svmb200_params.accum_chunk
STOP_NAMES = {0: "converged", 1: "not_found", 2: "infeasible_uv", 3: "nonpositive_eta", 4: "max_iter"}


class Params(C.Structure):
    _fields_ = [("C", C.c_double), ("gamma", C.c_double), ("eps", C.c_double), ("tau", C.c_double),
                ("max_iter", C.c_int), ("sv_tol", C.c_double), ("wss", C.c_int)]


class Result(C.Structure):
    _fields_ = [("num_iter", C.c_int), ("stop_reason", C.c_int), ("i_high", C.c_int64),
                ("i_low", C.c_int64), ("b_high", C.c_double), ("b_low", C.c_double),
                ("rows_computed", C.c_int64)]


def build(force: bool = False) -> None:
    """Compile the checker (building it is not using it)."""
    if force or not os.path.exists(os.path.join(HERE, "liboracle.so")) or \
            os.path.getmtime(os.path.join(HERE, "liboracle.so")) < os.path.getmtime(os.path.join(HERE, "smo_oracle.c")):
        subprocess.check_call(["make", "-s", "-C", HERE, "liboracle.so"])
    if os.path.exists("/root/reference/code/main3.cpp"):
        ref = os.path.join(HERE, "_ref", "libref_main3.so")
        if force or not os.path.exists(ref) or os.path.getmtime(ref) < os.path.getmtime(os.path.join(HERE, "ref_wrapper.cpp")):
            subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


_P = np.ctypeslib.ndpointer
_f64 = lambda: _P(dtype=np.float64, flags="C_CONTIGUOUS")
_f32 = lambda: _P(dtype=np.float32, flags="C_CONTIGUOUS")
_i32 = lambda: _P(dtype=np.int32, flags="C_CONTIGUOUS")
_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(os.path.join(HERE, "liboracle.so"))
        L.oracle_default_params.argtypes = [C.POINTER(Params)]
        L.oracle_kernel.restype = C.c_double
        L.oracle_kernel.argtypes = [_f64(), _f64(), C.c_size_t, C.c_double]
        L.oracle_find_min_max.argtypes = [_f64(), _f64(), _f64(), C.c_size_t, C.c_size_t]
        L.oracle_scale_features.argtypes = [_f64(), _f64(), _f64(), C.c_size_t, C.c_size_t]
        L.oracle_calc_i_high.restype = C.c_size_t
        L.oracle_calc_i_high.argtypes = [_f64(), _i32(), _f64(), C.c_size_t, C.c_double, C.c_double]
        L.oracle_calc_i_low.restype = C.c_size_t
        L.oracle_calc_i_low.argtypes = [_f64(), _i32(), _f64(), C.c_size_t, C.c_double, C.c_double]
        L.oracle_calculate_U_V.restype = C.c_int
        L.oracle_calculate_U_V.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double,
                                           C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.oracle_smo_train.restype = C.c_int
        L.oracle_smo_train.argtypes = [_f64(), _i32(), C.c_size_t, C.c_size_t, _f64(), C.POINTER(C.c_double),
                                       C.POINTER(Params), C.c_int, C.c_int, C.POINTER(Result),
                                       C.c_void_p, C.c_size_t, C.c_void_p]
        L.oracle_sv_indices.restype = C.c_size_t
        L.oracle_sv_indices.argtypes = [_f64(), C.c_size_t, C.c_double, C.c_void_p]
        L.oracle_predict.restype = C.c_int
        L.oracle_predict.argtypes = [_f64(), _i32(), C.c_size_t, _f64(), _i32(), _f64(), _i32(), C.c_size_t,
                                     C.c_size_t, C.c_double, C.c_double, C.c_int, C.c_int, C.c_void_p]
        L.oracle_device_exp.restype = C.c_double
        L.oracle_device_exp.argtypes = [C.c_double]
        L.oracle_device_dot.restype = C.c_double
        L.oracle_device_dot.argtypes = [_f32(), _f32(), C.c_size_t]
        L.oracle_device_sqdist.restype = C.c_double
        L.oracle_device_sqdist.argtypes = [_f32(), _f32(), C.c_size_t]
        L.oracle_set_device_chunk.argtypes = [C.c_int]
        L.oracle_device_kernel_f64.restype = C.c_double
        L.oracle_device_kernel_f64.argtypes = [_f64(), _f64(), C.c_size_t, C.c_double]
        L.oracle_device_kernel.restype = C.c_double
        L.oracle_device_kernel.argtypes = [_f32(), _f32(), C.c_size_t, C.c_double]
        L.oracle_u8_bind.argtypes = [_P(dtype=np.uint8, flags="C_CONTIGUOUS"), C.c_size_t, C.c_size_t]
        L.oracle_u8_layout.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.oracle_u8_sqdist.restype = C.c_double
        L.oracle_u8_sqdist.argtypes = [C.c_size_t, C.c_size_t]
        L.oracle_u8_kernel.restype = C.c_double
        L.oracle_u8_kernel.argtypes = [C.c_size_t, C.c_size_t, C.c_double]
        _lib = L
    return _lib


def have_ref() -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libref_main3.so"))


def ref():
    """The unmodified reference code/main3.cpp (oracle/_ref); None when it was never built."""
    global _ref
    if _ref is None:
        build()
        if not have_ref():
            return None
        R = C.CDLL(os.path.join(HERE, "_ref", "libref_main3.so"))
        R.ref_kernel.restype = C.c_double
        R.ref_kernel.argtypes = [_f64(), _f64(), C.c_size_t]
        R.ref_find_min_max.argtypes = [_f64(), _f64(), _f64(), C.c_size_t, C.c_size_t]
        R.ref_scale_features.argtypes = [_f64(), _f64(), _f64(), C.c_size_t, C.c_size_t]
        R.ref_smo_train.argtypes = [_f64(), _i32(), C.c_size_t, C.c_size_t, _f64(), C.POINTER(C.c_double),
                                    C.c_double, C.POINTER(C.c_int)]
        R.ref_sv_indices.restype = C.c_size_t
        R.ref_sv_indices.argtypes = [_f64(), C.c_size_t, C.c_double, _i32()]
        R.ref_predict.restype = C.c_int
        R.ref_predict.argtypes = [_f64(), _i32(), C.c_size_t, _f64(), _i32(), _f64(), C.c_size_t, C.c_size_t,
                                  C.c_double, C.c_double]
        R.ref_smo_iterations.restype = C.c_int
        R.ref_smo_iterations.argtypes = [_f64(), _i32(), C.c_size_t, C.c_size_t, C.c_double, C.c_int,
                                         C.c_void_p, C.POINTER(C.c_int64)]
        _ref = R
    return _ref


def set_device_chunk(ch: int) -> None:
    """Select which svmb200_params.accum_chunk the ORACLE_ARITH_DEVICE arithmetic restates."""
    lib().oracle_set_device_chunk(ch)


def u8_bind(rows_u8: np.ndarray):
    """Bind the raw uint8 training rows for ARITH_DEVICE_U8 (then smo_train(scale(rows)[0], y, arith=ARITH_DEVICE_U8)).
    Returns (R, M, T): range of the main class, features in it, features in the tail."""
    rows = np.ascontiguousarray(rows_u8, dtype=np.uint8)
    lib().oracle_u8_bind(rows, rows.shape[0], rows.shape[1])
    R, M, T = C.c_int(), C.c_int64(), C.c_int64()
    lib().oracle_u8_layout(C.byref(R), C.byref(M), C.byref(T))
    return R.value, M.value, T.value


def u8_kernel(row: int, piv: int, gamma: float = 0.00125) -> float:
    return lib().oracle_u8_kernel(row, piv, gamma)


def default_params(**kw) -> Params:
    p = Params()
    lib().oracle_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def scale(x_raw: np.ndarray, min_val=None, max_val=None):
    """find_min_max + scale_features (code/main3.cpp:338-339); returns (scaled f64, min, max)."""
    x = np.ascontiguousarray(x_raw, dtype=np.float64).copy()
    n, d = x.shape
    if min_val is None:
        min_val = np.zeros(d); max_val = np.zeros(d)
        lib().oracle_find_min_max(x, min_val, max_val, n, d)
    lib().oracle_scale_features(x, np.ascontiguousarray(min_val), np.ascontiguousarray(max_val), n, d)
    return x, min_val, max_val


def smo_train(x: np.ndarray, y: np.ndarray, params: Params | None = None, alpha0: np.ndarray | None = None,
              arith: int = ARITH_F64, trace_cap: int = 0, want_f: bool = False):
    """oracle_smo_train.  alpha0 given -> warm start (init=false).  Returns dict."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.int32)
    n, d = x.shape
    p = params or default_params()
    alpha = np.zeros(n) if alpha0 is None else np.ascontiguousarray(alpha0, dtype=np.float64).copy()
    b = C.c_double(0.0)
    res = Result()
    trace = np.zeros((max(trace_cap, 1), 2), dtype=np.int32)
    f = np.zeros(max(n, 1))
    lib().oracle_smo_train(x, y, n, d, alpha, C.byref(b), C.byref(p), 1 if alpha0 is None else 0, arith,
                           C.byref(res), trace.ctypes.data if trace_cap else None, trace_cap,
                           f.ctypes.data if want_f else None)
    updates = res.num_iter - 1
    return {"alpha": alpha, "b": b.value, "num_iter": res.num_iter, "stop_reason": res.stop_reason,
            "i_high": res.i_high, "i_low": res.i_low, "b_high": res.b_high, "b_low": res.b_low,
            "rows_computed": res.rows_computed, "trace": trace[:min(trace_cap, updates)],
            "f": f[:n] if want_f else None}


def sv_indices(alpha: np.ndarray, tol: float = 1e-8) -> np.ndarray:
    alpha = np.ascontiguousarray(alpha, dtype=np.float64)
    idx = np.zeros(max(alpha.size, 1), dtype=np.int32)
    c = lib().oracle_sv_indices(alpha, alpha.size, tol, idx.ctypes.data)
    return idx[:c].copy()


def predict(x_test, y_test, x_train, y_train, alpha, b, gamma=0.00125, tol=1e-8, ge_zero=False,
            arith: int = ARITH_F64):
    x_test = np.ascontiguousarray(x_test, dtype=np.float64)
    x_train = np.ascontiguousarray(x_train, dtype=np.float64)
    y_test = np.ascontiguousarray(y_test, dtype=np.int32)
    y_train = np.ascontiguousarray(y_train, dtype=np.int32)
    alpha = np.ascontiguousarray(alpha, dtype=np.float64)
    sv = sv_indices(alpha, tol)
    m, d = x_test.shape
    scores = np.zeros(max(m, 1))
    correct = lib().oracle_predict(x_test, y_test, m, x_train, y_train, alpha,
                                   sv if sv.size else np.zeros(1, np.int32), sv.size, d, gamma, b,
                                   1 if ge_zero else 0, arith, scores.ctypes.data)
    return correct, scores[:m]
