"""Small end-to-end run for compute-sanitizer (one tool per gpurun call): upload, train (cache off / small cache), predict."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
X, dg = pkg.datagen.make_dataset(400); y = pkg.datagen.binary_labels(dg)
Xt, dgt = pkg.datagen.make_dataset(100, split=1); yt = pkg.datagen.binary_labels(dgt)
s = pkg.Solver(0)
s.set_train_u8(X, y)
a = s.smo_train(cache_rows=0, max_iter=120)
b = s.smo_train(cache_rows=8, max_iter=120)
assert np.array_equal(a["alpha"], b["alpha"])
c, sc = s.predict_u8(Xt, yt)
w = s.smo_train(alpha0=a["alpha"], max_iter=20)
e = pkg.Solver(0); e.set_exact(True); e.set_train_u8(X, y); ee = e.smo_train(max_iter=60, cache_rows=4)
print("ok", a["num_iter"], b["cache_hits"], c, w["num_iter"], ee["num_iter"])
