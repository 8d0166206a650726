"""One-vs-rest over all ten digits on one upload of X (60 000 x 784): svmb200_set_labels + keep_cache."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
X, dg = pkg.datagen.make_dataset(n)
s = pkg.Solver(0)
res = {"n": n}
for label, keep in (("fresh_cache_per_digit", 0), ("kept_cache", 1)):
    s.set_train_u8(X, pkg.datagen.binary_labels(dg, 0))
    s.smo_train(max_iter=10, cache_rows=-1)
    s.set_train_u8(X, pkg.datagen.binary_labels(dg, 0))       # drops the cache
    t0 = time.perf_counter(); per = []
    for digit in range(10):
        s.set_labels(pkg.datagen.binary_labels(dg, digit))
        o = s.smo_train(cache_rows=-1, keep_cache=keep)
        per.append({"digit": digit, "iterations": o["num_iter"] - 1, "x_passes": o["x_passes"], "n_sv": o["n_sv"], "smo_ms": round(o["smo_ms"], 1)})
    res[label] = {"secs_10_digits": round(time.perf_counter() - t0, 3), "x_passes": sum(p["x_passes"] for p in per),
                  "iterations": sum(p["iterations"] for p in per), "per_digit": per}
print(json.dumps(res))
