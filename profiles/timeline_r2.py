"""Per-phase clock64 timeline of CTA 0 (a cluster leader) with the cluster exchange (profiling build
libsvmb200_timing.so; never loaded by the product).   python profiles/timeline_r2.py N ITERS CACHE_ROWS [storage]
Stamps of exchange g: 3 warp winners in shared memory, 4 after the CTA barrier, 8 the partner CTA's warp winners have arrived,
5 the cluster's winners folded, 6 record stored, 11 warp 0's share of the records seen, 12 every warp's share staged (barrier), 13 all
records folded, 7 exchange done;
of the iteration that follows: 0 loop top, 15 pivot rows arrived, 1 scalar update done, 2 pass done."""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
import svmb200.binding as B
B.LIB_PATH = os.path.join(B.PKG_DIR, "libsvmb200_timing.so")
n = int(sys.argv[1]); k = int(sys.argv[2]); cache = int(sys.argv[3]) if len(sys.argv) > 3 else 0
storage = sys.argv[4] if len(sys.argv) > 4 else None
X, dg = pkg.datagen.make_dataset(n); y = pkg.datagen.binary_labels(dg)
s = pkg.Solver(0)
if storage:
    s.set_storage(storage)
s.set_train_u8(X, y)
o = s.smo_train(max_iter=k, cache_rows=cache)
rows = min(4000, o["num_iter"] - 2)
t = np.zeros((rows, 16), dtype=np.int64)
assert s.L.svmb200_debug_timing(t.ctypes.data_as(C.c_void_p), rows) == 0
t = t[20:rows - 1]
nxt = np.roll(t, -1, axis=0)[:-1]; t = t[:-1]
ghz = 1.965
print(f"n={n} storage={storage} iters={o['num_iter'] - 1} x_passes={o['x_passes']} cache_hits={o['cache_hits']} "
      f"us/iter={1e3 * o['smo_ms'] / (o['num_iter'] - 1):.2f} grid={o['grid_blocks']}")
seg = {"pivot rows requested->arrived (0->15)": t[:, 15] - t[:, 0], "K12 + scalar update (15->1)": t[:, 1] - t[:, 15],
       "K12 + scalar update of a cached iteration (0->1)": t[:, 1] - t[:, 0], "pass (1->2)": t[:, 2] - t[:, 1], "warp winners (2->3')": nxt[:, 3] - t[:, 2], "CTA barrier (3'->4')": nxt[:, 4] - nxt[:, 3],
       "partner's warp winners there (4'->8')": nxt[:, 8] - nxt[:, 4], "fold of the cluster's warps (8'->5')": nxt[:, 5] - nxt[:, 8],
       "record store (5'->6')": nxt[:, 6] - nxt[:, 5],
       "poll: warp 0's records seen (6'->11')": nxt[:, 11] - nxt[:, 6], "barrier: all warps' records (11'->12')": nxt[:, 12] - nxt[:, 11],
       "fold of all records (12'->13')": nxt[:, 13] - nxt[:, 12], "to partner, done (13'->7')": nxt[:, 7] - nxt[:, 13],
       "to loop top (7'->0')": nxt[:, 0] - nxt[:, 7], "exchange (2->0')": nxt[:, 0] - t[:, 2], "iteration (0->0')": nxt[:, 0] - t[:, 2] + t[:, 2] - t[:, 0]}
passlen = t[:, 2] - t[:, 1]
cached = np.zeros(len(t), bool) if cache == 0 else passlen < np.median(passlen[passlen > 0]) * 0.6 if o["cache_hits"] < o["x_passes"] else np.ones(len(t), bool)
for label, mask in (("X-pass iterations", ~cached), ("cached iterations", cached)):
    if mask.sum() < 5:
        continue
    print(f"{label} ({mask.sum()} in the sample)")
    for name, v in seg.items():
        if ("(0->1)" in name) != (label == "cached iterations") and ("15" in name or "(0->1)" in name):
            continue
        v = v[mask]; v = v[(v > 0) & (v < 10 ** 9)]
        if len(v):
            print(f"  {name:42s} median {np.median(v) / ghz / 1e3:7.3f} us   p90 {np.percentile(v, 90) / ghz / 1e3:7.3f} us")
if k > 110:
    G = o["grid_blocks"]
    g = np.zeros(8 * 160 * 4, dtype=np.int64)
    assert s.L.svmb200_debug_timing_grid(g.ctypes.data_as(C.c_void_p)) == 0
    g = g[:8 * G * 4].reshape(8, G, 4)
    for it in (2, 3, 4):
        t0 = g[it, :, 0].min()
        top, sc, pe = (g[it, :, 0] - t0) / 1e3, (g[it, :, 1] - t0) / 1e3, (g[it, :, 2] - t0) / 1e3
        nxt_top = (g[it + 1, :, 0] - t0) / 1e3
        print(f"  iter {100 + it}: loop-top spread {top.max():.2f} us | pass end min/median/max {pe.min():.2f}/{np.median(pe):.2f}/{pe.max():.2f} us"
              f" | pass len min/med/max {(pe - sc).min():.2f}/{np.median(pe - sc):.2f}/{(pe - sc).max():.2f} | next loop top {nxt_top.min():.2f}..{nxt_top.max():.2f}")
