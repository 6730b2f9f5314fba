#!/usr/bin/env python3
"""Turns an ncu report into the short text summary kept under profiles/ (the .ncu-rep itself is scratch).
usage: ncu_summary.py report.ncu-rep > profiles/ncu_<what>_rNN_summary.txt
Per captured launch: duration, warp instructions, issue / occupancy, DRAM and L2 bytes, the largest stall
reasons, shared-memory bank conflicts; then the source lines that carry most instructions (-lineinfo)."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, launches = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
WANT = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__grid_size"]
for n, r in enumerate(launches):
    print("== launch %d: %s" % (n, r[col["Kernel Name"]][:110]))
    for w in WANT:
        if w in col:
            print("   %-62s %s %s" % (w, r[col[w]], units[col[w]]))
    stalls = [(float(r[i] or 0), h.split("issue_stalled_")[1].split("_per_issue")[0]) for h, i in col.items()
              if "smsp__average_warps_issue_stalled_" in h and h.endswith("_per_issue_active.ratio") and r[i]]
    print("   stalls per issue: " + ", ".join("%s %.2f" % (nm, v) for v, nm in sorted(stalls, reverse=True)[:6]))
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--launch-skip",
                          str(n), "--launch-count", "1"], capture_output=True, text=True).stdout
    cur, agg = None, {}
    for s in csv.reader(io.StringIO(src)):
        if len(s) == 2 and s[0] == "File Path":
            cur = s[1].split("/")[-1]
            continue
        if len(s) < 8 or s[0] == "Line No" or s[2] != "-":
            continue
        try:
            agg[(cur, int(s[0]))] = (int(s[7]), int(s[6]), s[1].strip()[:84])
        except ValueError:
            pass
    ti = sum(v[0] for v in agg.values()) or 1
    ts = sum(v[1] for v in agg.values()) or 1
    print("   hottest source lines (share of warp instructions / of stall samples):")
    for (f, ln), (i, sm, text) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:10]:
        print("     %-16s %4d %5.1f%% %5.1f%% | %s" % (f, ln, 100 * i / ti, 100 * sm / ts, text))
