"""Small end-to-end run for compute-sanitizer: upload, train (cache off / small cache, first- and second-order selection),
predict, in the three storages, plus rows whose tails are long enough for 16-row tail blocks with several blocks per warp."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
X, dg = pkg.datagen.make_dataset(6000); y = pkg.datagen.binary_labels(dg)
Xt, dgt = pkg.datagen.make_dataset(100, split=1); yt = pkg.datagen.binary_labels(dgt)
s = pkg.Solver(0)
s.set_train_u8(X, y)                                  # uint8 storage: 40 rows per CTA, a block of 5 rows per warp
a = s.smo_train(cache_rows=0, max_iter=60)
b = s.smo_train(cache_rows=8, max_iter=60)
assert np.array_equal(a["alpha"], b["alpha"])
c, sc = s.predict_u8(Xt, yt)
w = s.smo_train(alpha0=a["alpha"], max_iter=10)
g = s.smo_train(cache_rows=8, max_iter=20, wss=2)
f = pkg.Solver(0); f.set_storage("f32"); f.set_train_u8(X[:600], y[:600]); ff = f.smo_train(max_iter=40, cache_rows=4)
e = pkg.Solver(0); e.set_storage("f64"); e.set_train_u8(X[:600], y[:600]); ee = e.smo_train(max_iter=40, cache_rows=4)
rng = np.random.default_rng(5)
d, n = 832, 24000                                     # no shared range: all tail, 16-row tail blocks, 20 rows per warp
ranges = np.asarray([1 + k % 254 for k in range(d)])
L = (rng.integers(0, 256 - ranges, size=d) + rng.integers(0, ranges + 1, size=(n, d))).astype(np.uint8)
yl = np.where(rng.random(n) < 0.5, 1, -1).astype(np.int32)
t = pkg.Solver(0); t.set_train_u8(L, yl); tt = t.smo_train(max_iter=3, cache_rows=0, gamma=0.5 / d, C=1.0)
print("ok", a["num_iter"], b["cache_hits"], c, w["num_iter"], g["num_iter"], ff["num_iter"], ee["num_iter"], tt["num_iter"])
