"""Fixed cost of one persistent-kernel launch inside the CUDA-event bracket: kernel time against the number of iterations
(60 000 x 784, uint8 rows, row cache off); intercept = launch + prologue (weights, barriers, ring fill, selection over the initial
state) + epilogue, slope = one iteration."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
X, dg = pkg.datagen.make_dataset(n); y = pkg.datagen.binary_labels(dg)
s = pkg.Solver(0)
s.set_train_u8(X, y)
for _ in range(5):
    s.smo_train(max_iter=50, cache_rows=0)
ks = [1, 2, 3, 5, 10, 20, 50, 100, 300]
med = []
for k in ks:
    t = [s.smo_train(max_iter=k, cache_rows=0)["smo_ms"] * 1e3 for _ in range(15)]
    med.append(float(np.median(t)))
    print(f"max_iter={k:4d}  kernel {med[-1]:8.2f} us  min {min(t):8.2f}")
A = np.vstack([np.ones(len(ks)), ks]).T
c = np.linalg.lstsq(A, np.asarray(med), rcond=None)[0]
print(f"fit: {c[0]:.2f} us + {c[1]:.3f} us/iteration")
