#!/usr/bin/env python3
"""Per-source-line instruction and stall-sample shares from
`ncu -i X.ncu-rep --page source --print-source cuda,sass --csv > src.csv` (kernels built with
-lineinfo, captured with --import-source on).  usage: hot_lines.py src.csv [top N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, agg = None, {}
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) < 8 or r[0] == "Line No" or r[2] != "-":      # SASS rows carry an address in column 2
        continue
    try:
        ln, inst, samp = int(r[0]), int(r[7]), int(r[6])
    except ValueError:
        continue
    agg[(cur, ln)] = (inst, samp, r[1].strip()[:90])
tot = sum(v[0] for v in agg.values()) or 1
ts = sum(v[1] for v in agg.values()) or 1
print("warp instructions", tot, "samples", ts)
for (f, ln), (i, s, src) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-14s %4d %5.1f%% inst %5.1f%% samples | %s" % (f, ln, 100 * i / tot, 100 * s / ts, src))
