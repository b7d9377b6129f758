#!/usr/bin/env python
"""Rank CUDA source lines of an `ncu --page source --csv --print-source cuda,sass` dump by executed instructions.
usage: ncu_lines.py <src.csv> [<raw.csv>]"""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
cur, hdr, out = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1]
    elif len(r) > 5 and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0] != "":
        d = dict(zip(hdr, r))
        try:
            out.append((cur.split("/")[-1], int(r[0]), int(d["Instructions Executed"]), int(d["# Samples"]), r[1][:100]))
        except ValueError:
            pass
tot, ts = sum(o[2] for o in out) or 1, sum(o[3] for o in out) or 1
print("total warp instructions %d, samples %d" % (tot, ts))
for o in sorted(out, key=lambda x: -x[2])[:int(sys.argv[3]) if len(sys.argv) > 3 else 40]:
    print("%-20s L%4d inst %5.1f%%  smp %5.1f%%  %s" % (o[0], o[1], 100.0 * o[2] / tot, 100.0 * o[3] / ts, o[4]))
if len(sys.argv) > 2 and sys.argv[2] != "-":
    rows = list(csv.reader(open(sys.argv[2])))
    h, u, d = rows[0], rows[1], rows[2]
    for w in ("gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
              "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__registers_per_thread",
              "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "smsp__thread_inst_executed_per_inst_executed.ratio",
              "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"):
        if w in h:
            print(w, "=", d[h.index(w)], u[h.index(w)])
    for i, x in enumerate(h):
        if "issue_stalled" in x and "per_issue_active" in x:
            print("stall", x.split("stalled_")[1].split("_per")[0], d[i])
