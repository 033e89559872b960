#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page) per kernel: time, DRAM bytes, throughputs, occupancy, stall mix."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(out.splitlines())); h = r[0]
def col(n): return h.index(n) if n in h else -1
M = [("ms", "gpu__time_duration.sum"), ("GB_rd", "dram__bytes_read.sum"), ("GB_wr", "dram__bytes_write.sum"),
     ("dram%", "dram__throughput.avg.pct_of_peak_sustained_elapsed"), ("l1%", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
     ("l2%", "lts__throughput.avg.pct_of_peak_sustained_elapsed"), ("fp64%", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed"),
     ("issue%", "sm__issue_active.avg.pct_of_peak_sustained_elapsed"), ("occ%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
     ("regs", "launch__registers_per_thread"),
     ("st_long", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
     ("st_short", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"),
     ("st_bar", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"),
     ("st_wait", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"),
     ("st_math", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"),
     ("st_mio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"),
     ("st_lg", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"),
     ("inst_M", "smsp__inst_executed.sum")]
units = r[1]
print("%-44s" % "kernel" + "".join("%9s" % m[0] for m in M))
for row in r[2:]:
    name = row[col("Kernel Name")].replace("void mhh::<unnamed>::", "")[:43]
    vals = []
    for lab, m in M:
        i = col(m)
        if i < 0: vals.append("      n/a"); continue
        v = float(row[i].replace(",", "")); u = units[i]
        if lab == "ms": v = v / 1e6 if u == "ns" else (v / 1e3 if u == "us" else v)
        if lab.startswith("GB"): v = v / {"byte": 1e9, "Kbyte": 1e6, "Mbyte": 1e3, "Gbyte": 1.}.get(u, 1e9)
        if lab == "inst_M": v = v / 1e6
        vals.append("%9.3f" % v)
    print("%-44s" % name + "".join(vals))
