#!/usr/bin/env python3
"""SASS instruction mix of the hot kernels of libdmgraph.so (cuobjdump -sass): the 14 most frequent
mnemonics per kernel plus the counts that answer "is anything vectorised / asynchronous":
LDG/STG widths, shared-memory atomics, remote-capable atomics (ATOMG/RED), votes and shuffles.
usage: python profiles/sass_mix.py > profiles/sass_mix_rNN.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOT = ["khop_msbfs_kernel", "khop_emit_bits_kernel", "khop_large_kernel", "khop1_warp_kernel", "khop_global_kernel",
       "search_multi_kernelILi0", "search_emit_log_kernel", "closure_p2p_kernel", "adjacency_kernel",
       "adjacency_long_kernel", "qx_fill_kernel", "ord_level_kernelILi1"]
out = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "domain-mcp-server_b200", "libdmgraph.so")],
                     capture_output=True, text=True).stdout
cur, mix = None, {}
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        mix[cur] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        mix[cur][m.group(2)] += 1
seen = set()
for name in sorted(mix):
    key = next((h for h in HOT if h in name), None)
    if not key or (key, name.count("ILi")) in seen and key not in ("khop_msbfs_kernel",):
        continue
    if key in seen:
        continue
    seen.add(key)
    c = mix[name]
    total = sum(c.values())
    print("== %s  (%d instructions)" % (name, total))
    for op, n in c.most_common(14):
        print("   %6d %s" % (n, op))
    pick = lambda pre: {k: v for k, v in c.items() if k.startswith(pre)}
    notable = {}
    for pre in ("LDG.E.128", "LDG.E.64", "STG.E.128", "STG.E.64", "LDS.128", "STS.128", "ATOMS", "ATOMG", "RED", "VOTE",
                "SHFL", "MATCH", "LDGSTS", "UBLKCP", "UTMA", "MEMBAR", "NANOSLEEP"):
        n = sum(pick(pre).values())
        if n:
            notable[pre] = n
    print("   notable: " + ", ".join("%s x%d" % kv for kv in notable.items()))
