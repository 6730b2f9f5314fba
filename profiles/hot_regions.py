#!/usr/bin/env python3
"""Summarise an `ncu --page source --csv` dump: contiguous SASS regions with equal execution
counts (loop bodies), their share of executed instructions and of stall samples.
usage: hot_regions.py src.csv [kernel-index]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
idx = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
k = int(sys.argv[2]) if len(sys.argv) > 2 else 0
sec = rows[idx[k] + 2: idx[k + 1] if k + 1 < len(idx) else None]
print(rows[idx[k]][1][:80])
tot = sum(int(r[5]) for r in sec)
samp = sum(int(r[4]) for r in sec)
print('total warp instructions', tot, 'samples', samp)
cur, blocks = None, []
for i, r in enumerate(sec):
    n, s = int(r[5]), int(r[4])
    if cur and cur['n'] == n:
        cur['cnt'] += 1
        cur['samp'] += s
        cur['end'] = i
    else:
        cur = {'n': n, 'cnt': 1, 'samp': s, 'start': i, 'end': i}
        blocks.append(cur)
for b in blocks:
    if b['n'] * b['cnt'] > tot * 0.01 or b['samp'] > samp * 0.02:
        print('lines %4d-%4d: exec=%11d x %3d instrs = %5.1f%% inst, %5.1f%% samples | %s' % (
            b['start'], b['end'], b['n'], b['cnt'], 100 * b['n'] * b['cnt'] / tot, 100 * b['samp'] / samp,
            sec[b['start']][1].strip()[:50]))
