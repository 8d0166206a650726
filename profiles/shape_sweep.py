"""us per iteration of the persistent kernel at other row lengths (random rows, row cache off)."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
shapes = [(400000, 30), (200000, 100), (120000, 257), (100000, 500), (60000, 600), (40000, 1000)]
s = pkg.Solver(0)
for n, d in shapes:
    rng = np.random.default_rng(d)
    X = rng.integers(0, 256, (n, d), dtype=np.uint8)
    y = np.where(rng.random(n) < 0.3, 1, -1).astype(np.int32)
    s.set_train_u8(X, y)
    s.smo_train(max_iter=30, cache_rows=0, gamma=0.5 / d)
    ms = []
    for _ in range(3):
        o = s.smo_train(max_iter=500, cache_rows=0, gamma=0.5 / d)
        ms.append(o["smo_ms"] / (o["num_iter"] - 1))
    us = 1e3 * float(np.median(ms))
    row_bytes = ((d + 3) // 4) * 16
    print(json.dumps({"n": n, "d": d, "us_per_iter": round(us, 2), "alg_GBps": round(n * row_bytes / us / 1e3, 1), "threads_env": os.environ.get("SVMB200_THREADS", "default")}), flush=True)
