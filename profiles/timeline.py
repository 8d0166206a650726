"""Per-phase clock64 timeline of CTA 0 (profiling build libsvmb200_timing.so; never used by the product).
Stamps per exchange g: 3 after warp reduce, 4 after CTA barrier, 5 after record store, 6 after poll, 7 after merge;
then of the iteration that follows: 0 loop top, 1 scalar update done (pivots loaded, K12, alpha), 2 pass done."""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
import svmb200.binding as B
B.LIB_PATH = os.path.join(B.PKG_DIR, "libsvmb200_timing.so")
n = int(sys.argv[1]); k = int(sys.argv[2]); cache = int(sys.argv[3]) if len(sys.argv) > 3 else 0
X, dg = pkg.datagen.make_dataset(n); y = pkg.datagen.binary_labels(dg)
s = pkg.Solver(0)
s.set_train_u8(X, y)
o = s.smo_train(max_iter=k, cache_rows=cache)
rows = min(k, 4000)
t = np.zeros((rows, 16), dtype=np.int64)
assert s.L.svmb200_debug_timing(t.ctypes.data_as(C.c_void_p), rows) == 0
t = t[20:rows - 1]                     # skip the cold start
nxt = np.roll(t, -1, axis=0)[:-1]; t = t[:-1]
ghz = 1.965
seg = {"  pivot loads issued->arrived (0->15)": t[:, 15] - t[:, 0], "  K12+alpha math (15->1)": t[:, 1] - t[:, 15],
       "  pub: smem read (4'->8')": nxt[:, 8] - nxt[:, 4], "  pub: warp reduce (8'->9')": nxt[:, 9] - nxt[:, 8],
       "  pub: record stores (9'->10')": nxt[:, 10] - nxt[:, 9], "  pub: red (10'->5')": nxt[:, 5] - nxt[:, 10],
       "  rd: barrier (6'->11')": nxt[:, 11] - nxt[:, 6], "  rd: record read (11'->12')": nxt[:, 12] - nxt[:, 11],
       "  rd: warp reduce (12'->13')": nxt[:, 13] - nxt[:, 12], "  rd: barrier (13'->14')": nxt[:, 14] - nxt[:, 13],
       "  rd: merge (14'->7')": nxt[:, 7] - nxt[:, 14],
       "scalar (0->1)": t[:, 1] - t[:, 0], "pass (1->2)": t[:, 2] - t[:, 1], "warp reduce (2->3')": nxt[:, 3] - t[:, 2],
       "CTA barrier (3'->4')": nxt[:, 4] - nxt[:, 3], "CTA reduce+publish (4'->5')": nxt[:, 5] - nxt[:, 4], "spin on counter (5'->6')": nxt[:, 6] - nxt[:, 5],
       "read records+merge (6'->7')": nxt[:, 7] - nxt[:, 6], "to loop top (7'->0')": nxt[:, 0] - nxt[:, 7], "iteration (0->0')": nxt[:, 0] - t[:, 0]}
print(f"n={n} iters={o['num_iter'] - 1} x_passes={o['x_passes']} cache_hits={o['cache_hits']} us/iter={1e3 * o['smo_ms'] / (o['num_iter'] - 1):.2f}")
for name, v in seg.items():
    v = v[(v > 0) & (v < 10 ** 9)]
    print(f"  {name:38s} median {np.median(v) / ghz / 1e3:7.3f} us   p90 {np.percentile(v, 90) / ghz / 1e3:7.3f} us")
# per-CTA globaltimer stamps of iterations 100..107: 0 loop top, 1 scalar done, 2 pass done
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
    # is the arrival skew systematic?  per-CTA pass-end offset (relative to the iteration's median), iterations 101..106
    offs = np.array([(g[it, :, 2] - np.median(g[it, :, 2])) / 1e3 for it in range(1, 7)])
    print("  per-CTA pass-end offset: mean over iterations min/max %.2f/%.2f us, std across iterations (median CTA) %.2f us" %
          (offs.mean(0).min(), offs.mean(0).max(), np.median(offs.std(0))))
    c = np.corrcoef(offs)
    print("  correlation of the offset vector between iterations: mean off-diagonal %.2f" % ((c.sum() - np.trace(c)) / (c.size - len(c))))
    late = np.argsort(-offs.mean(0))[:8]
    print("  latest CTAs:", [(int(b), round(float(offs.mean(0)[b]), 2)) for b in late])
