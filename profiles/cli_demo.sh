# svm_train on idx files, 60 000 x 784 train / 10 000 test (synthetic), the reference's printed lines
set -e
python - <<PY
import sys
sys.path.insert(0, "."); import __graft_entry__ as e
pkg = e.load_package()
X, dg = pkg.datagen.make_dataset(60000); Xt, dgt = pkg.datagen.make_dataset(10000, split=1)
pkg.datagen.write_idx("/tmp/tri", "/tmp/trl", X, dg); pkg.datagen.write_idx("/tmp/tei", "/tmp/tel", Xt, dgt)
PY
B=parallelizing-support-vector-machine-training-with-gpu-and-mpi_b200/svm_train
for opts in "" "--exact" "--tensor-cores" "--cache-rows 0"; do echo "== svm_train $opts"; $B --train-images /tmp/tri --train-labels /tmp/trl --test-images /tmp/tei --test-labels /tmp/tel $opts; done
