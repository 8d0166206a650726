"""Decision function over 10 000 test rows for the 60k model: exact CUDA-core kernel vs tensor cores (tcgen05 TF32x3 + recheck)."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
pkg = entry.load_package()
s = pkg.Solver(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
X, dg = pkg.datagen.make_dataset(n); y = pkg.datagen.binary_labels(dg)
Xt, dgt = pkg.datagen.make_dataset(10000, split=1); yt = pkg.datagen.binary_labels(dgt)
res = {"n": n, "m": 10000}
for storage in ("u8", "f32"):
    s.set_storage(storage)
    s.set_train_u8(X, y)
    o = s.smo_train(cache_rows=-1)
    r = {"n_sv": o["n_sv"]}
    for label, tc in (("exact_cuda_cores", False),) + ((("tensor_cores_tf32x3", True),) if storage == "f32" else ()):
        s.set_tensor_cores(tc, 0.05)
        s.predict_u8(Xt[:256], yt[:256])
        ts = []
        for _ in range(5):
            t0 = time.perf_counter(); c, sc = s.predict_u8(Xt, yt); ts.append(time.perf_counter() - t0)
        r[label] = {"secs_median_incl_upload": round(float(np.median(ts)), 5), "correct": c, "rechecked": s.tc_rechecked() if tc else None}
        if tc:
            r[label]["max_abs_score_error"] = float(np.max(np.abs(sc - ref_scores)))
        else:
            ref_scores = sc
    s.set_tensor_cores(False)
    res[storage + "_storage"] = r
print(json.dumps(res))
