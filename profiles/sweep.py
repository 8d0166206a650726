"""Experiment driver (not the bench): iterations/s of the persistent kernel over knobs and sizes."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry

pkg = entry.load_package()
peak = 6549.4
sizes = [int(a) for a in sys.argv[1].split(",")] if len(sys.argv) > 1 else [60000]
chunks = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 2, 4]
persists = [int(a) for a in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0, 1]
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 2000
s = pkg.Solver(0)
Xall, dg = pkg.datagen.make_dataset(max(sizes))
yall = pkg.datagen.binary_labels(dg)
for n in sizes:
    s.set_train_u8(Xall[:n], yall[:n])
    for ch in chunks:
        for lp in persists:
            s.smo_train(max_iter=50, cache_rows=0, accum_chunk=ch, l2_persist=lp)
            ms = []
            for _ in range(3):
                o = s.smo_train(max_iter=steps, cache_rows=0, accum_chunk=ch, l2_persist=lp)
                ms.append(o["smo_ms"] / (o["num_iter"] - 1))
            us = 1e3 * float(np.median(ms))
            gbs = n * 3136 / us / 1e3
            print(json.dumps({"n": n, "chunk": ch, "l2_persist": lp, "us_per_iter": round(us, 2), "iters": o["num_iter"] - 1,
                              "alg_GBps": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 3)}), flush=True)
