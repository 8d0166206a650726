"""Turns what the GPU box brought back in gpurun_out/ into the tracked summaries of this directory (runs without a GPU):
  python profiles/ncu_summarize.py rep  gpurun_out/ncu_r2_smo60k_u8.ncu-rep  profiles/ncu_r2_smo60k_u8.txt  u8  300  "<command line>"
      one `ncu --set full` capture -> the metrics DESIGN.md / r2_notes.md quote, and this storage's entry of profiles/traffic_r2.json
  python profiles/ncu_summarize.py launches  gpurun_out/launches_r2.csv  profiles/launches_r2_summary.csv  "<command line>"
      the launch list (gpu__time_duration.sum per launch) -> launches, total time and share per kernel
"""
import csv
import datetime
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__cluster_size",
        "launch__grid_size", "launch__block_size", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__shared_mem_per_block_dynamic"]
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def head():
    return subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()


def rep(path, out, storage, iters, cmd):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    names, units, vals = rows[0], rows[1], rows[2]
    col = {n: i for i, n in enumerate(names)}
    kernel = vals[col["Kernel Name"]]
    lines = [f"ncu --set full --clock-control none --import-source on -k regex:smo_persistent -c 1: {cmd}  (HEAD {head()}; {iters} iterations in one "
             f"launch, row cache off)", f"kernel: {kernel}"]
    got = {}
    for m in WANT:
        if m in col:
            lines.append(f"{m} [{units[col[m]]}] = {vals[col[m]]}")
            got[m] = (float(vals[col[m]].replace(",", "")), units[col[m]])
    lines.append("note: ncu reports launch__cluster_* = 0 for this cudaLaunchKernelEx launch (cooperative + cluster dimension attributes); the "
                 "library's own record of the launch (svmb200_stats.cluster_size, printed by bench.py) is 2, also when run under ncu, and "
                 "cudaOccupancyMaxActiveClusters returns 74 clusters of 2 for the 148-CTA grid")
    open(out, "w").write("\n".join(lines) + "\n")
    dram = sum(got[m][0] * SCALE[got[m][1]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    ms = got["gpu__time_duration.sum"][0] * {"ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}[got["gpu__time_duration.sum"][1]]
    tpath = os.path.join(ROOT, "profiles", "traffic_r2.json")
    t = json.load(open(tpath)) if os.path.exists(tpath) else {}
    t[storage] = {"kernel": kernel, "file": os.path.relpath(out, ROOT), "captured": datetime.date.today().isoformat(), "head": head(),
                  "iterations_in_launch": iters, "dram_bytes_per_launch": dram, "dram_bytes_per_iteration": dram / iters,
                  "algorithmic_bytes_per_iteration": 188160000, "us_per_iter_under_ncu": 1e3 * ms / iters,
                  "l2_hit_rate_pct": got["lts__t_sector_hit_rate.pct"][0]}
    json.dump(t, open(tpath, "w"), indent=1)
    print(open(out).read())


def launches(path, out, cmd):
    text = open(path).read()
    start = text.index('"ID"')
    rows = list(csv.DictReader(io.StringIO(text[start:])))
    agg = {}
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[r["Metric Unit"]]
        a = agg.setdefault(r["Kernel Name"], [0, 0.0])
        a[0] += 1
        a[1] += v
    total = sum(v[1] for v in agg.values())
    with open(out, "w") as f:
        f.write(f"ncu --metrics gpu__time_duration.sum --clock-control none: {cmd} (N=1; HEAD {head()}); serialised, cold-cache: compare SHARES\n")
        f.write("kernel, launches, total_ms, share\n")
        for k, (c, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k}, {c}, {ms:.3f}, {100 * ms / total:.2f}%\n")
    print(open(out).read())


if __name__ == "__main__":
    if sys.argv[1] == "rep":
        rep(sys.argv[2], sys.argv[3], sys.argv[4], int(sys.argv[5]), sys.argv[6])
    else:
        launches(sys.argv[2], sys.argv[3], sys.argv[4])
