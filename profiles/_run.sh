set -x
python bench.py --steps 2000 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err || { tail -5 gpurun_out/bench_n1.err; exit 1; }
python bench.py --steps 300 --warmup 3 --no-cpu > gpurun_out/bench_small.json 2> gpurun_out/bench_small.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 300 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1
python profiles/profile_target.py 60000 300 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:smo_persistent -c 1 -o gpurun_out/prof_r1_smo60k_v2 python profiles/profile_target.py 60000 300 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
python bench.py --impl reference --steps 2000 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
