"""Where the uint8 pass of CTA 0's warp 0 spends its cycles (profiling build libsvmb200_timing.so): per iteration the sums of
wait-for-data, main-class sums, refill (fence + bulk copy issue), integer reduce, row loads / loop, tail walk, finishing."""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
import svmb200.binding as B
B.LIB_PATH = os.path.join(B.PKG_DIR, "libsvmb200_timing.so")
n = int(sys.argv[1]); k = int(sys.argv[2])
X, dg = pkg.datagen.make_dataset(n); y = pkg.datagen.binary_labels(dg)
s = pkg.Solver(0)
s.set_train_u8(X, y)
o = s.smo_train(max_iter=k, cache_rows=0)
rows = min(4000, o["num_iter"] - 2)
t = np.zeros((rows, 8), dtype=np.int64)
assert s.L.svmb200_debug_timing_pass(t.ctypes.data_as(C.c_void_p), rows) == 0
t = t[20:]
names = ["wait for the group's bytes", "main-class sums (LDS + IDP4A)", "wait for the tail block's bytes", "fence + refill issue", "integer warp reduce",
         "block setup (row state loads, loop)", "tail walk (a lane per row) + refill issue", "finishing (exp, f update, selection)"]
print(f"n={n} us/iter={1e3 * o['smo_ms'] / (o['num_iter'] - 1):.2f}; cycles of warp 0 of CTA 0 per iteration (median), 1.965 GHz")
tot = 0
for i, nm in enumerate(names):
    v = np.median(t[:, i]); tot += v
    print(f"  {nm:40s} {v:9.0f} cyc = {v / 1965:6.2f} us")
print(f"  {'sum':40s} {tot:9.0f} cyc = {tot / 1965:6.2f} us")
