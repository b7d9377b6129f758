#!/usr/bin/env python
"""SASS instructions of given source lines with their executed counts, from `ncu --page source --csv --print-source cuda,sass`.
usage: ncu_line_sass.py <src.csv> <file-suffix> <line> [<line> ...]"""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
want = set(int(x) for x in sys.argv[3:])
cur, line = None, None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1]
    elif len(r) > 8 and r[0] not in ("", "Line No"):
        try:
            line = int(r[0])
        except ValueError:
            line = None
        if cur and cur.endswith(sys.argv[2]) and line in want:
            print("L%d: %s   [%s executed]" % (line, r[1].strip()[:110], r[7]))
    elif len(r) > 8 and r[0] == "" and cur and cur.endswith(sys.argv[2]) and line in want:
        print("      %-60s %s" % (r[3].strip(), r[7]))
