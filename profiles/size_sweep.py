"""BASELINE.json configs[2]: training-size sweep 10k..60k x 784 on 1 B200 (the shape of the reference's
gpu_svm_main4 loop, code/gpu_svm4.sh:16-21), row cache off / on, full training to convergence through the C ABI
from host buffers; plus the batched decision function over 10k test rows."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
s = pkg.Solver(0)
Xall, dg = pkg.datagen.make_dataset(60000); yall = pkg.datagen.binary_labels(dg)
Xt, dgt = pkg.datagen.make_dataset(10000, split=1); yt = pkg.datagen.binary_labels(dgt)
rows = []
for n in (10000, 20000, 30000, 40000, 50000, 60000):
    rec = {"n": n}
    for label, cache in (("cache_off", 0), ("cache_on", -1)):
        s.set_train_u8(Xall[:n], yall[:n]); s.smo_train(max_iter=20, cache_rows=cache)      # warm-up
        t0 = time.perf_counter()
        s.set_train_u8(Xall[:n], yall[:n])
        o = s.smo_train(cache_rows=cache)
        secs = time.perf_counter() - t0
        rec[label] = {"train_secs": round(secs, 4), "smo_secs": round(o["smo_ms"] / 1e3, 4), "iterations": o["num_iter"] - 1,
                      "us_per_iter": round(1e3 * o["smo_ms"] / (o["num_iter"] - 1), 2), "x_passes": o["x_passes"], "n_sv": o["n_sv"]}
    s.predict_u8(Xt[:100], yt[:100])
    t0 = time.perf_counter()
    correct, _ = s.predict_u8(Xt, yt)
    rec["predict_10k_secs"] = round(time.perf_counter() - t0, 4)
    rec["test_correct"] = correct
    g = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", f"golden_n{n}.npz")
    if os.path.exists(g):
        gg = np.load(g)
        rec["sv_set_equals_reference"] = bool(np.array_equal(s.sv_ids(), gg["sv_ids"]))
        rec["test_correct_reference"] = int(gg["correct"])
    rows.append(rec)
    print(json.dumps(rec), flush=True)
json.dump(rows, open("gpurun_out/size_sweep.json", "w"), indent=1)
