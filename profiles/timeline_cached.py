"""Per-phase clock64 timeline of CTA 0 for CACHED iterations only (profiling build libsvmb200_timing.so)."""
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
o = s.smo_train(max_iter=k, cache_rows=-1)
rows = min(4000, o["num_iter"] - 2)
t = np.zeros((rows, 16), dtype=np.int64)
assert s.L.svmb200_debug_timing(t.ctypes.data_as(C.c_void_p), rows) == 0
t = t[20:rows - 1]
nxt = np.roll(t, -1, axis=0)[:-1]; t = t[:-1]
ghz = 1.965
cached = (t[:, 2] - t[:, 1]) < 6 * ghz * 1e3          # pass shorter than 6 us
print(f"n={n} iters={o['num_iter'] - 1} x_passes={o['x_passes']} cache_hits={o['cache_hits']} us/iter={1e3 * o['smo_ms'] / (o['num_iter'] - 1):.2f}; cached iterations in the sample: {cached.sum()} of {len(cached)}")
seg = {"scalar (0->1)": t[:, 1] - t[:, 0], "pass (1->2)": t[:, 2] - t[:, 1], "warp reduce (2->3')": nxt[:, 3] - t[:, 2],
       "CTA barrier (3'->4')": nxt[:, 4] - nxt[:, 3], "CTA reduce+publish (4'->5')": nxt[:, 5] - nxt[:, 4], "spin on counter (5'->6')": nxt[:, 6] - nxt[:, 5],
       "  rd: barrier (6'->11')": nxt[:, 11] - nxt[:, 6], "  rd: record read (11'->12')": nxt[:, 12] - nxt[:, 11],
       "  rd: warp reduce (12'->13')": nxt[:, 13] - nxt[:, 12], "  rd: barrier (13'->14')": nxt[:, 14] - nxt[:, 13],
       "  rd: merge (14'->7')": nxt[:, 7] - nxt[:, 14],
       "read records+merge (6'->7')": nxt[:, 7] - nxt[:, 6], "to loop top (7'->0')": nxt[:, 0] - nxt[:, 7], "iteration (0->0')": nxt[:, 0] - t[:, 0]}
for label, mask in (("cached", cached), ("X pass", ~cached)):
    print(label)
    for name, v in seg.items():
        v = v[mask]; v = v[(v > 0) & (v < 10 ** 9)]
        if len(v): print(f"  {name:38s} median {np.median(v) / ghz / 1e3:7.3f} us   p90 {np.percentile(v, 90) / ghz / 1e3:7.3f} us")
