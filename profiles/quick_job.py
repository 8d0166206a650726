"""Whole training jobs at 60k (cache off / on / exact path), SMO time only: a quick A/B for kernel changes."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
X, dg = pkg.datagen.make_dataset(n); y = pkg.datagen.binary_labels(dg)
out = {}
s = pkg.Solver(0); s.set_train_u8(X, y)
s.smo_train(max_iter=50, cache_rows=0)
for label, cr in (("cache_off", 0), ("cache_on", -1)):
    ms = []
    for _ in range(3):
        o = s.smo_train(cache_rows=cr); ms.append(o["smo_ms"])
    out[label] = {"smo_ms": round(float(np.median(ms)), 2), "iters": o["num_iter"] - 1, "x_passes": o["x_passes"], "n_sv": o["n_sv"], "b": o["b"]}
s.close()
if "--exact" in sys.argv:
    s = pkg.Solver(0); s.set_exact(True); s.set_train_u8(X, y)
    s.smo_train(max_iter=50, cache_rows=0)
    o = s.smo_train(max_iter=1000, cache_rows=0); out["exact_us_per_iter_cache_off"] = round(1e3 * o["smo_ms"] / (o["num_iter"] - 1), 2)
    o = s.smo_train(cache_rows=-1); out["exact_cache_on"] = {"smo_ms": round(o["smo_ms"], 2), "iters": o["num_iter"] - 1, "n_sv": o["n_sv"]}
print(json.dumps(out))
