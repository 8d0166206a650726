"""Repeatability stress: the same problem solved over and over (cache off / on / tiny, fp32 / exact storage) must give the
same bits every time -- a race in the record exchange or the ring would show up as a difference or a hang."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
Xall, dg = pkg.datagen.make_dataset(20000); yall = pkg.datagen.binary_labels(dg)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
t0 = time.time(); runs = 0
for exact in (False, True):
    s = pkg.Solver(0); s.set_exact(exact)
    for n in (777, 5000, 20000):
        s.set_train_u8(Xall[:n], yall[:n])
        ref = s.smo_train(cache_rows=0, trace_cap=100000); tr = s.get_trace(100000)
        for r in range(reps):
            for cache in (0, -1, 7):
                o = s.smo_train(cache_rows=cache, trace_cap=100000)
                assert np.array_equal(o["alpha"], ref["alpha"]) and o["b"] == ref["b"] and o["num_iter"] == ref["num_iter"], (exact, n, r, cache)
                assert np.array_equal(s.get_trace(100000), tr), (exact, n, r, cache)
                runs += 1
        print(f"exact={exact} n={n}: {reps * 3} repeats identical ({ref['num_iter'] - 1} iterations each)", flush=True)
    s.close()
print(f"stress ok: {runs} training runs in {time.time() - t0:.1f} s")
