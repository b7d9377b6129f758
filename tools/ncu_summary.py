#!/usr/bin/env python
"""Transpose an `ncu -i X.ncu-rep --page raw --csv` dump into metric rows x kernel columns, keeping the metrics
the roofline discussion uses, and (optionally) refresh profiles/ncu_traffic.json for a workload.
usage: ncu_summary.py <raw.csv> <out_summary.csv> [<workload> <reads> <source text>]"""
import csv
import json
import os
import re
import sys

KEEP = re.compile(r"^(dram__bytes|gpu__time_duration|l1tex__t_(requests|sectors)_pipe_lsu_mem_global_op_ld\.sum|launch__(block_size|grid_size|occupancy_limit|registers)"
                  r"|lts__t_sector_hit_rate|l1tex__t_sector_hit_rate|sm__inst_executed_pipe_(alu|lsu)|sm__maximum_warps|sm__throughput|sm__warps_active"
                  r"|smsp__average_warps_issue_stalled_.*_per_issue_active|smsp__inst_executed\.sum$|smsp__thread_inst_executed_per_inst_executed"
                  r"|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$|launch__shared_mem_per_block)")
rows = list(csv.reader(open(sys.argv[1])))
h, u, data = rows[0], rows[1], rows[2:]
names = [re.sub(r"^void ", "", re.sub(r"\(.*", "", d[h.index("Kernel Name")])) for d in data]
with open(sys.argv[2], "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["metric", "unit"] + names)
    for i, m in enumerate(h):
        if KEEP.match(m):
            w.writerow([m, u[i]] + [d[i] for d in data])
print("kernels:", names)
if len(sys.argv) > 3:
    wl, reads, src = sys.argv[3], int(sys.argv[4]), sys.argv[5]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "ncu_traffic.json")
    t = json.load(open(path))
    entry = {"reads": reads, "source": src}
    ir, iw = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
    group = {"k_seed": "k_seed", "k_cand": "k_cand", "k_verify": "k_verify", "k_compact": "k_compact", "k_trace": "k_trace", "k_sam": "k_sam"}
    for d, n in zip(data, names):
        for prefix, key in group.items():
            if n.startswith(prefix):
                entry[key] = entry.get(key, 0) + float(d[ir]) * scale[u[ir]] + float(d[iw]) * scale[u[iw]]
    t[wl] = entry
    json.dump(t, open(path, "w"), indent=1)
    print(wl, entry)
