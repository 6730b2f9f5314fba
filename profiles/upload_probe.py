import sys, time, os
sys.path.insert(0, "domain-mcp-server_b200"); sys.path.insert(0, ".")
import numpy as np, dmgraph
from dmgraph import synth
os.environ["DMG_UPLOAD_TIMING"] = "1"
eng = dmgraph.Engine(0)
sg = synth.big_graph_fast(4_000_000, 100_000_000, cap=2000, entry_frac=0.01, seed=synth.SEED_C5)
print("dtypes", sg.src.dtype, sg.dst.dtype, sg.src.flags["C_CONTIGUOUS"], os.cpu_count(), flush=True)
for _ in range(2):
    t0 = time.perf_counter()
    g = dmgraph.Graph(eng, sg.n_vertices, sg.src, sg.dst, sg.entry)
    print("upload_s", time.perf_counter() - t0, flush=True)
    g.release()
