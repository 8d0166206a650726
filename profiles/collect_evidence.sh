#!/bin/bash
# What a round's single-GPU evidence run executes on the GPU box (gpurun -- bash profiles/collect_evidence.sh): smoke, the GPU tests, the
# bench lines, the timelines, then -- only after the plain commands exited 0 -- the two ncu captures and the launch list.  The results
# land in gpurun_out/; profiles/ncu_summarize.py and profiles/sass_counts.py turn them into the tracked summaries.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
O=gpurun_out
S=$O/summary_ev.txt
: > $S
timeout -k 10 300 python -c "import __graft_entry__ as e; e.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $S
timeout -k 10 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $S; tail -3 $O/pytest.log >> $S
timeout -k 10 900 python bench.py --steps 2000 --warmup 3 > $O/bench_r2_n1.json 2> $O/bench_r2_n1.err; echo "bench 2000 rc=$?" >> $S
timeout -k 10 900 python bench.py --steps 20 --warmup 3 > $O/bench_r2_n1_steps20.json 2> $O/bench_20.err; echo "bench 20 rc=$?" >> $S
timeout -k 10 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_r2_reference_arm.json 2> $O/bench_ref.err; echo "bench ref rc=$?" >> $S
for st in u8 f32; do
  timeout -k 10 300 python profiles/timeline_r2.py 60000 600 0 $st > $O/timeline_r2_60k_$st.txt 2>&1
  timeout -k 10 300 python profiles/timeline_r2.py 60000 3000 -1 $st > $O/timeline_r2_cached_$st.txt 2>&1
done
timeout -k 10 300 python profiles/timeline_pass.py 60000 600 > $O/timeline_r2_pass_u8.txt 2>&1
echo "timelines done" >> $S
# ncu only after the plain commands exited 0
for st in u8 f32; do
  timeout -k 10 300 python profiles/profile_target.py 60000 300 0 -1 $st > $O/plain_$st.log 2>&1 && \
  timeout -k 10 900 ncu --set full --clock-control none --import-source on -k regex:smo_persistent -c 1 -f -o $O/ncu_r2_smo60k_$st python profiles/profile_target.py 60000 300 0 -1 $st > $O/ncu_$st.log 2>&1
  echo "ncu $st rc=$?" >> $S
done
timeout -k 10 600 python bench.py --steps 300 --warmup 3 --no-cpu --no-exact --no-sweep --no-cascade > $O/plain_bench.log 2>&1 && \
timeout -k 10 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_r2.csv python bench.py --steps 300 --warmup 3 --no-cpu --no-exact --no-sweep --no-cascade > $O/ncu_launch.log 2>&1
echo "launch list rc=$?" >> $S
cat $S
