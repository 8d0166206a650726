"""The command profiled by ncu: one persistent-kernel launch of K iterations at N x 784, row cache off."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 300
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 0
l2p = int(sys.argv[4]) if len(sys.argv) > 4 else 0
storage = sys.argv[5] if len(sys.argv) > 5 else "f32"
X, dg = pkg.datagen.make_dataset(n)
y = pkg.datagen.binary_labels(dg)
s = pkg.Solver(0)
s.set_storage(storage)
s.set_train_u8(X, y)
o = s.smo_train(max_iter=k, cache_rows=0, accum_chunk=chunk, l2_persist=l2p)
print({kk: o[kk] for kk in ("num_iter", "smo_ms", "launches", "grid_blocks")}, "us/iter", 1e3 * o["smo_ms"] / (o["num_iter"] - 1))
