#!/usr/bin/env python3
"""Small driver for profiling khop_global: k-hop OUT on a 1M/30M graph (bench_configs c4 shape).
usage: khop3_probe.py [k] [n_queries] [scratch GiB, comma separated; 0 = library default]"""
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "domain-mcp-server_b200"))
import dmgraph
from dmgraph import synth
k = int(sys.argv[1]) if len(sys.argv) > 1 else 3
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
budgets = [float(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0.0]
sg = synth.big_graph(1_000_000, 30_000_000, cap=500, entry_frac=0.01, seed=synth.SEED_C4)
rng = np.random.Generator(np.random.PCG64(synth.SEED_C4 + 1))
ids = rng.integers(0, sg.n_vertices, nq).astype(np.uint32)
q_off = np.arange(nq + 1, dtype=np.uint32)
for b in budgets:
    eng = dmgraph.Engine(0, scratch_bytes=int(b * (1 << 30)))
    g = dmgraph.Graph(eng, sg.n_vertices, sg.src, sg.dst, sg.entry)
    eng.profile(True)
    for _ in range(2):
        eng.profile_reset()
        off, out, hop = g.neighbors_batch(dmgraph.OUT, k, 0, q_off, ids)
    print("scratch GiB", b, eng.last_stats())
    print({k_: (int(v[0]), round(v[1] / v[0], 3)) for k_, v in eng.profile_table().items()}, flush=True)
    g.release()
    eng.close()
