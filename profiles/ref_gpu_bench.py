"""The reference's own CUDA program (code/gpu_svm_main3.cu, unmodified, built for sm_100a by oracle/Makefile `refgpu`)
timed on this box on the same synthetic data as bench.py: "reference CUDA on B200".  It reads mnist3_{train,test}_data.csv
from its working directory, so the CSVs are written to a scratch directory first."""
import json, os, re, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry
pkg = entry.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
limit = int(sys.argv[2]) if len(sys.argv) > 2 else 900
exe = os.path.join(ROOT, "oracle", "_ref", "gpu_svm3")
work = os.path.join(ROOT, "gpurun_out", f"refgpu_{n}")
os.makedirs(work, exist_ok=True)
X, dg = pkg.datagen.make_dataset(n); Xt, dgt = pkg.datagen.make_dataset(10000, split=1)
t0 = time.time()
pkg.datagen.write_csv(os.path.join(work, "mnist3_train_data.csv"), X, dg)
pkg.datagen.write_csv(os.path.join(work, "mnist3_test_data.csv"), Xt, dgt)
t_csv = time.time() - t0
t0 = time.time()
try:
    r = subprocess.run([exe], cwd=work, capture_output=True, text=True, timeout=limit)
    out, rc = r.stdout + r.stderr, r.returncode
except subprocess.TimeoutExpired as e:
    out, rc = (e.stdout or b"").decode() + (e.stderr or b"").decode(), "timeout"
wall = time.time() - t0
open(os.path.join(ROOT, "gpurun_out", f"refgpu_{n}.log"), "w").write(out)
for f in ("mnist3_train_data.csv", "mnist3_test_data.csv"):
    os.remove(os.path.join(work, f))
lines = [l for l in out.splitlines() if re.search(r"time|iter|accuracy|support|SV|b =|n =", l, re.I)]
print(json.dumps({"n": n, "rc": rc, "wall_secs_incl_csv_read": round(wall, 1), "csv_write_secs": round(t_csv, 1), "lines": lines[-14:]}))
