"""ctypes binding of include/svmb200.h (the drop-in boundary of the solver path)."""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
LIB_PATH = os.path.join(PKG_DIR, "libsvmb200.so")
HEADER = os.path.join(ROOT, "include", "svmb200.h")
STOP_NAMES = {0: "converged", 1: "not_found", 2: "infeasible_uv", 3: "nonpositive_eta", 4: "max_iter", 5: "peer_lost"}


class SvmB200Error(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [("C", C.c_double), ("gamma", C.c_double), ("eps", C.c_double), ("tau", C.c_double),
                ("max_iter", C.c_int32), ("sv_tol", C.c_double), ("wss", C.c_int32), ("warm_start", C.c_int32),
                ("cache_rows", C.c_int32), ("accum_chunk", C.c_int32), ("trace_cap", C.c_int32),
                ("l2_persist", C.c_int32), ("keep_cache", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("num_iter", C.c_int32), ("stop_reason", C.c_int32), ("i_high", C.c_int64), ("i_low", C.c_int64),
                ("b_high", C.c_double), ("b_low", C.c_double), ("smo_ms", C.c_double), ("init_ms", C.c_double),
                ("x_passes", C.c_int64), ("cache_hits", C.c_int64), ("x_bytes_streamed", C.c_int64),
                ("n_sv", C.c_int32), ("grid_blocks", C.c_int32), ("launches", C.c_int32), ("cluster_size", C.c_int32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class CascadeStats(C.Structure):
    _fields_ = [("rounds", C.c_int32), ("converged", C.c_int32), ("n_sv", C.c_int32), ("b", C.c_double), ("train_ms", C.c_double),
                ("smo_iterations", C.c_int64), ("stages", C.c_int32), ("solve_ms", C.c_double), ("smo_ms", C.c_double)]


def declared_symbols() -> list[str]:
    """Every function include/svmb200.h declares (used by the export test)."""
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(svmb200_[a-z0-9_]+)\s*\(", text)))


def build_library(force: bool = False) -> str:
    """nvcc -gencode arch=compute_100a,code=sm_100a ... -> libsvmb200.so, in tree."""
    srcs = [os.path.join(PKG_DIR, "csrc", f) for f in os.listdir(os.path.join(PKG_DIR, "csrc"))] + [HEADER]
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(s) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", PKG_DIR, "libsvmb200.so"])
    return LIB_PATH


_lib = None


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SvmB200Error(f"{LIB_PATH} is not built (run __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    L.svmb200_version.restype = C.c_char_p
    L.svmb200_default_params.argtypes = [C.POINTER(Params)]
    L.svmb200_default_params.restype = None
    L.svmb200_create.argtypes = [C.POINTER(vp), C.c_int]
    L.svmb200_destroy.argtypes = [vp]
    L.svmb200_destroy.restype = None
    L.svmb200_last_error.argtypes = [vp]
    L.svmb200_last_error.restype = C.c_char_p
    L.svmb200_set_train_u8.argtypes = [vp, vp, vp, i64, i32, vp, vp]
    L.svmb200_set_train_f64.argtypes = [vp, vp, vp, i64, i32, vp, vp]
    L.svmb200_set_train_scaled_f64.argtypes = [vp, vp, vp, i64, i32]
    L.svmb200_set_exact.argtypes = [vp, i32]
    L.svmb200_set_storage.argtypes = [vp, i32]
    L.svmb200_get_u8_layout.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
    L.svmb200_set_labels.argtypes = [vp, vp]
    L.svmb200_minmax_f64.argtypes = [vp, vp, i64, i32, vp, vp]
    L.svmb200_scale_f64.argtypes = [vp, vp, vp, vp, i64, i32]
    L.svmb200_smo_train.argtypes = [vp, C.POINTER(Params), vp, C.POINTER(dbl), C.POINTER(Stats)]
    L.svmb200_sv_ids.argtypes = [vp, vp, i32, C.POINTER(i32)]
    for name in ("svmb200_predict_u8", "svmb200_predict_f64", "svmb200_predict_scaled_f64"):
        getattr(L, name).argtypes = [vp, vp, vp, i64, i32, C.POINTER(i32), vp]
    L.svmb200_get_trace.argtypes = [vp, vp, i32, C.POINTER(i32)]
    L.svmb200_get_f.argtypes = [vp, vp]
    L.svmb200_kernel_row.argtypes = [vp, i64, dbl, i32, vp]
    L.svmb200_select_pair.argtypes = [vp, vp, vp, vp, i64, dbl, dbl, C.POINTER(i64), C.POINTER(i64)]
    L.svmb200_get_x_scaled.argtypes = [vp, vp]
    L.svmb200_pair_update.argtypes = [vp, i32, i32, dbl, dbl, dbl, dbl, dbl, dbl, dbl, vp, C.POINTER(i32)]
    L.svmb200_set_tensor_cores.argtypes = [vp, i32, dbl]
    L.svmb200_debug_tc_dots.argtypes = [vp, vp, C.POINTER(i64)]
    L.svmb200_save_model.argtypes = [vp, C.c_char_p]
    L.svmb200_load_model.argtypes = [vp, C.c_char_p]
    L.svmb200_shard_setup.argtypes = [vp, i32, i32, vp]
    L.svmb200_shard_connect.argtypes = [vp, i32, vp]
    L.svmb200_set_model.argtypes = [vp, vp, dbl]
    L.svmb200_minmax_u8_noalloc.argtypes = [C.c_int, vp, i64, i32, vp, vp, vp]
    L.svmb200_cascade_unique_id.argtypes = [vp]
    L.svmb200_cascade_connect.argtypes = [vp, i32, i32, vp]
    L.svmb200_cascade_set_partition_u8.argtypes = [vp, vp, vp, i64, i32, i64, i64]
    L.svmb200_cascade_train.argtypes = [vp, i32, C.POINTER(Params), i32, C.POINTER(CascadeStats)]
    L.svmb200_cascade_get_model.argtypes = [vp, vp, vp, i32, C.POINTER(i32)]
    _lib = L
    return L


def default_params(**kw) -> Params:
    p = Params()
    load_library().svmb200_default_params(C.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise TypeError(f"unknown parameter {k}")
        setattr(p, k, v)
    return p


def cascade_unique_id() -> bytes:
    """128 bytes that every rank of a cascade passes to Solver.cascade_connect (made on one rank, handed to the others)."""
    L = load_library()
    buf = C.create_string_buffer(128)
    rc = L.svmb200_cascade_unique_id(buf)
    if rc != 0:
        raise SvmB200Error(f"svmb200_cascade_unique_id = {rc}: {L.svmb200_last_error(None).decode()}")
    return buf.raw


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Solver:
    """One context on one GPU.  Mirrors the reference's call sequence (code/main3.cpp:338-405):
    set_train (find_min_max + scale_features) -> smo_train (SMO_train) -> sv_ids (get_SV_indices)
    -> predict."""

    def __init__(self, device: int = 0):
        self.L = load_library()
        h = C.c_void_p()
        rc = self.L.svmb200_create(C.byref(h), device)
        if rc != 0:
            raise SvmB200Error(f"svmb200_create({device}) = {rc}: {self.L.svmb200_last_error(None).decode()}")
        self.h = h
        self.n = 0
        self.d = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.svmb200_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc, what):
        if rc != 0:
            raise SvmB200Error(f"{what} = {rc}: {self.L.svmb200_last_error(self.h).decode()}")

    # ---- ingest
    def set_labels(self, y):
        y = np.ascontiguousarray(y, dtype=np.int32)
        assert y.size == self.n
        self._ck(self.L.svmb200_set_labels(self.h, _ptr(y)), "set_labels")

    def set_exact(self, enable: bool):
        """FP64 storage of X for the next set_train_* (the exact path); False = fp32 (default)."""
        self._ck(self.L.svmb200_set_exact(self.h, 1 if enable else 0), "set_exact")

    STORAGE = {"auto": -1, "f32": 0, "f64": 1, "u8": 2}

    def set_storage(self, storage: str):
        """How the next set_train_* keeps X: "auto" (default: bytes stay bytes, everything else fp32), "f32", "f64" (exact),
        "u8" (set_train_u8 only: the bytes themselves)."""
        self._ck(self.L.svmb200_set_storage(self.h, self.STORAGE[storage]), "set_storage")

    def u8_layout(self):
        r, m, t, b = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        self._ck(self.L.svmb200_get_u8_layout(self.h, C.byref(r), C.byref(m), C.byref(t), C.byref(b)), "get_u8_layout")
        return {"main_range": r.value, "main_features": m.value, "tail_features": t.value, "row_bytes": b.value}

    def set_train_u8(self, X, y):
        X = np.ascontiguousarray(X, dtype=np.uint8)
        y = np.ascontiguousarray(y, dtype=np.int32)
        n, d = X.shape
        mn, mx = np.zeros(d), np.zeros(d)
        self._ck(self.L.svmb200_set_train_u8(self.h, _ptr(X), _ptr(y), n, d, _ptr(mn), _ptr(mx)), "set_train_u8")
        self.n, self.d = n, d
        return mn, mx

    def set_train_f64(self, X, y):
        X = np.ascontiguousarray(X, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.int32)
        n, d = X.shape
        mn, mx = np.zeros(d), np.zeros(d)
        self._ck(self.L.svmb200_set_train_f64(self.h, _ptr(X), _ptr(y), n, d, _ptr(mn), _ptr(mx)), "set_train_f64")
        self.n, self.d = n, d
        return mn, mx

    def set_train_scaled(self, Xs, y):
        Xs = np.ascontiguousarray(Xs, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.int32)
        n, d = Xs.shape
        self._ck(self.L.svmb200_set_train_scaled_f64(self.h, _ptr(Xs), _ptr(y), n, d), "set_train_scaled_f64")
        self.n, self.d = n, d

    def minmax_f64(self, X):
        X = np.ascontiguousarray(X, dtype=np.float64)
        n, d = X.shape
        mn, mx = np.zeros(d), np.zeros(d)
        self._ck(self.L.svmb200_minmax_f64(self.h, _ptr(X), n, d, _ptr(mn), _ptr(mx)), "minmax_f64")
        return mn, mx

    def scale_f64(self, X, mn, mx):
        X = np.ascontiguousarray(X, dtype=np.float64).copy()
        n, d = X.shape
        mn = np.ascontiguousarray(mn, dtype=np.float64); mx = np.ascontiguousarray(mx, dtype=np.float64)
        self._ck(self.L.svmb200_scale_f64(self.h, _ptr(X), _ptr(mn), _ptr(mx), n, d), "scale_f64")
        return X

    # ---- solver
    def smo_train(self, params: Params | None = None, alpha0=None, **kw):
        p = params or default_params(**kw)
        if alpha0 is not None:
            p.warm_start = 1
            alpha = np.ascontiguousarray(alpha0, dtype=np.float64).copy()
        else:
            alpha = np.zeros(self.n)
        b = C.c_double()
        st = Stats()
        self._ck(self.L.svmb200_smo_train(self.h, C.byref(p), _ptr(alpha), C.byref(b), C.byref(st)), "smo_train")
        out = st.as_dict()
        out.update(alpha=alpha, b=b.value)
        return out

    def sv_ids(self):
        c = C.c_int32()
        self._ck(self.L.svmb200_sv_ids(self.h, None, 0, C.byref(c)), "sv_ids")
        ids = np.zeros(max(c.value, 1), dtype=np.int32)
        self._ck(self.L.svmb200_sv_ids(self.h, _ptr(ids), c.value, C.byref(c)), "sv_ids")
        return ids[:c.value]

    def _predict(self, fn, X, dtype, y_test, ge_zero):
        X = np.ascontiguousarray(X, dtype=dtype)
        m = X.shape[0]
        yt = None if y_test is None else np.ascontiguousarray(y_test, dtype=np.int32)
        scores = np.zeros(m)
        correct = C.c_int32()
        self._ck(fn(self.h, _ptr(X), _ptr(yt), m, 1 if ge_zero else 0, C.byref(correct), _ptr(scores)), "predict")
        return correct.value, scores

    def predict_u8(self, X, y_test=None, ge_zero=False):
        return self._predict(self.L.svmb200_predict_u8, X, np.uint8, y_test, ge_zero)

    def predict_f64(self, X, y_test=None, ge_zero=False):
        return self._predict(self.L.svmb200_predict_f64, X, np.float64, y_test, ge_zero)

    def predict_scaled(self, Xs, y_test=None, ge_zero=False):
        return self._predict(self.L.svmb200_predict_scaled_f64, Xs, np.float64, y_test, ge_zero)

    def set_tensor_cores(self, enable: bool, recheck_threshold: float = 0.05):
        self._ck(self.L.svmb200_set_tensor_cores(self.h, 1 if enable else 0, recheck_threshold), "set_tensor_cores")

    def tc_dots(self):
        dots = np.zeros((128, 128), dtype=np.float32)
        r = C.c_int64()
        self._ck(self.L.svmb200_debug_tc_dots(self.h, _ptr(dots), C.byref(r)), "debug_tc_dots")
        return dots

    def tc_rechecked(self) -> int:
        r = C.c_int64()
        self._ck(self.L.svmb200_debug_tc_dots(self.h, None, C.byref(r)), "debug_tc_dots")
        return r.value

    def save_model(self, path: str):
        self._ck(self.L.svmb200_save_model(self.h, path.encode()), "save_model")

    def load_model(self, path: str):
        self._ck(self.L.svmb200_load_model(self.h, path.encode()), "load_model")

    # ---- row sharding over several GPUs (one process per GPU)
    def shard_setup(self, rank: int, world: int) -> bytes:
        buf = C.create_string_buffer(64)
        self._ck(self.L.svmb200_shard_setup(self.h, rank, world, buf), "shard_setup")
        return buf.raw

    def shard_connect(self, world: int, handles: bytes):
        assert len(handles) == 64 * world
        self._ck(self.L.svmb200_shard_connect(self.h, world, C.c_char_p(handles)), "shard_connect")

    def set_model(self, alpha, b):
        alpha = np.ascontiguousarray(alpha, dtype=np.float64)
        self._ck(self.L.svmb200_set_model(self.h, _ptr(alpha), float(b)), "set_model")

    # ---- Cascade SVM over several GPUs (one rank per GPU)
    MODES = {"star": 0, "tree": 1}

    def cascade_connect(self, rank: int, world: int, unique_id: bytes):
        assert len(unique_id) == 128
        self._ck(self.L.svmb200_cascade_connect(self.h, rank, world, C.c_char_p(unique_id)), "cascade_connect")

    def cascade_set_partition_u8(self, X_part, y_part, first_id: int, n_total: int):
        X_part = np.ascontiguousarray(X_part, dtype=np.uint8)
        y_part = np.ascontiguousarray(y_part, dtype=np.int32)
        n, d = X_part.shape
        self._ck(self.L.svmb200_cascade_set_partition_u8(self.h, _ptr(X_part), _ptr(y_part), n, d, first_id, n_total), "cascade_set_partition_u8")
        self.d = d

    def cascade_train(self, mode: str = "star", max_rounds: int = 50, params: Params | None = None, **kw):
        p = params or default_params(**kw)
        st = CascadeStats()
        self._ck(self.L.svmb200_cascade_train(self.h, self.MODES[mode], C.byref(p), max_rounds, C.byref(st)), "cascade_train")
        out = {k: getattr(st, k) for k, _ in st._fields_}
        self.n = out["n_sv"]
        return out

    def cascade_get_model(self):
        c = C.c_int32()
        self._ck(self.L.svmb200_cascade_get_model(self.h, None, None, 0, C.byref(c)), "cascade_get_model")
        ids = np.zeros(max(c.value, 1), dtype=np.int32)
        alpha = np.zeros(max(c.value, 1))
        self._ck(self.L.svmb200_cascade_get_model(self.h, _ptr(ids), _ptr(alpha), c.value, C.byref(c)), "cascade_get_model")
        return ids[:c.value], alpha[:c.value]

    # ---- introspection
    def get_trace(self, cap):
        pairs = np.zeros((max(cap, 1), 2), dtype=np.int32)
        c = C.c_int32()
        self._ck(self.L.svmb200_get_trace(self.h, _ptr(pairs), cap, C.byref(c)), "get_trace")
        return pairs[:c.value]

    def get_f(self):
        f = np.zeros(self.n)
        self._ck(self.L.svmb200_get_f(self.h, _ptr(f)), "get_f")
        return f

    def kernel_row(self, i, gamma=0.00125, accum_chunk=0):
        row = np.zeros(self.n)
        self._ck(self.L.svmb200_kernel_row(self.h, i, gamma, accum_chunk, _ptr(row)), "kernel_row")
        return row

    def select_pair(self, alpha, f, y, C_=10.0, eps=1e-12):
        alpha = np.ascontiguousarray(alpha, dtype=np.float64)
        f = np.ascontiguousarray(f, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.int32)
        ih, il = C.c_int64(), C.c_int64()
        self._ck(self.L.svmb200_select_pair(self.h, _ptr(alpha), _ptr(f), _ptr(y), alpha.size, C_, eps,
                                            C.byref(ih), C.byref(il)), "select_pair")
        return ih.value, il.value

    def pair_update(self, y_high, y_low, a_high, a_low, b_high, b_low, K12, C_=10.0, eps=1e-12):
        """calculate_U_V + the pair update as the solver runs them -> (dict, stop)."""
        out = np.zeros(7)
        stop = C.c_int32()
        self._ck(self.L.svmb200_pair_update(self.h, y_high, y_low, a_high, a_low, b_high, b_low, K12, C_, eps, _ptr(out),
                                            C.byref(stop)), "pair_update")
        return dict(zip(("U", "V", "eta", "alpha_high", "alpha_low", "c_high", "c_low"), out.tolist())), stop.value

    def get_x_scaled(self):
        Xs = np.zeros((self.n, self.d), dtype=np.float32)
        self._ck(self.L.svmb200_get_x_scaled(self.h, _ptr(Xs)), "get_x_scaled")
        return Xs
