#!/usr/bin/env python3
"""dev helper: per-source-line stall samples / thread efficiency of one .ncu-rep (needs -lineinfo + --import-source on)"""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); h, u, v = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "launch__registers_per_thread", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__pcsamp_warps_issue_stalled_long_scoreboard", "smsp__pcsamp_warps_issue_stalled_no_instructions", "smsp__pcsamp_warps_issue_stalled_wait",
        "smsp__pcsamp_warps_issue_stalled_branch_resolving", "smsp__pcsamp_warps_issue_stalled_selected", "smsp__pcsamp_warps_issue_stalled_short_scoreboard",
        "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle", "smsp__pcsamp_warps_issue_stalled_lg_throttle", "smsp__pcsamp_warps_issue_stalled_not_selected",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"]
for i, k in enumerate(h):
    if k in want:
        print("%-70s %-12s %s" % (k, u[i], v[i]))
mix = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
cur = None; agg = {}; tot = 0
for r in csv.reader(mix.splitlines()):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if len(r) < 10 or r[0] in ("Line No", ""):
        continue
    try:
        ln = int(r[0]); s = int(r[4]); inst = int(r[7]); thr = int(r[8])
    except ValueError:
        continue
    agg[(cur, ln)] = (s, inst, thr, r[1][:100]); tot += s
byfile = {}
for (f, l), (s, i, t, src) in agg.items():
    b = byfile.setdefault(f, [0, 0, 0]); b[0] += s; b[1] += i; b[2] += t
print("total samples", tot)
for f, (s, i, t) in sorted(byfile.items(), key=lambda x: -x[1][0]):
    print("%-28s %5.1f%%  warp-inst %10d  avg thr %.1f" % (f, 100 * s / tot, i, t / max(i, 1)))
for (f, l), (s, i, t, src) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print("%5.1f%% %-12s %4d thr=%4.1f inst=%8d %s" % (100 * s / tot, f, l, t / max(i, 1), i, src.strip()))
