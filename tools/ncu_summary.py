#!/usr/bin/env python
"""Summarise an `ncu --set full` report (raw page CSV) into a small markdown table for profiles/.
usage: ncu -i rep.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_summary.py raw.csv > profiles/xxx.md"""
import csv
import sys

M = [("gpu__time_duration.sum", "time"),
     ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
     ("dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "dram active %"),
     ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64 pipe %"),
     ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
     ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
     ("launch__registers_per_thread", "regs"),
     ("smsp__inst_executed.sum", "warp inst"),
     ("lts__t_sector_hit_rate.pct", "L2 hit %"),
     ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_sb"),
     ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait"),
     ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_sb"),
     ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math_throttle"),
     ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall not_sel")]

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
print("| kernel | grid x block | " + " | ".join(n for _, n in M) + " |")
print("|---|---|" + "---|"*len(M))
for r in rows[2:]:
    cells = []
    for k, _ in M:
        if k not in idx:
            cells.append("n/a"); continue
        v, u = r[idx[k]], units[idx[k]]
        try:
            f = float(v.replace(",", ""))
            v = ("%.4g" % f)
        except ValueError:
            pass
        cells.append((v + " " + u).strip())
    name = r[idx["Kernel Name"]].replace("void ", "").split("(T1")[0]
    print("| `%s` | %s x %s | " % (name, r[idx["Grid Size"]], r[idx["Block Size"]]) + " | ".join(cells) + " |")
