#!/usr/bin/env python3
"""Multi-GPU through the C ABI alone (dmg_dist_*: one process, N devices, NCCL bound by the library):
run on a box with >= N GPUs, e.g.  gpurun --gpus 2 -- python tests/run_dist_capi.py --gpus 2
Checks, against the oracle: (1) a config-2-shaped query batch sharded over N replicas, (2) the
vertex-partitioned closure in both transports -- the persistent peer-memory kernel (default where the
devices have peer access) and the per-level NCCL allgather (DMG_DIST_NCCL=1) -- on the config-5-shaped
graph and on a sparse, deep one that stays on push levels.  One JSON line."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "domain-mcp-server_b200"))
sys.path.insert(0, ROOT)
import dmgraph  # noqa: E402
from dmgraph import synth  # noqa: E402
from oracle import oracle as ora  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=2)
ap.add_argument("--vertices", type=int, default=4_000_000)
ap.add_argument("--edges", type=int, default=100_000_000)
ap.add_argument("--queries", type=int, default=65536)
ap.add_argument("--steps", type=int, default=5)
args = ap.parse_args()
out = {"check": "dist_capi", "n_gpus": args.gpus}
dist = dmgraph.Dist(args.gpus)

# (1) replicated CSR, sharded query batch -- whole batch against the oracle
sg = synth.eco_graph(with_names=False)
q_off, q_src = synth.stacktrace_queries(sg, n_queries=args.queries)
dg = dmgraph.DistGraph(dist, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off)
off, ids, hop = dg.neighbors_batch(dmgraph.BOTH, 2, 1, q_off, q_src)
o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry, sg.part_off)
woff, wids, whop, _ = o.khop_batch(2, 2, 1, q_off, q_src, threads=ora.hardware_threads())
assert np.array_equal(off, woff) and np.array_equal(ids, wids) and np.array_equal(hop, whop)
bufs = (np.zeros(q_off.size, np.uint64), np.zeros(ids.size, np.uint32), np.zeros(ids.size, np.uint8))
ts = []
for _ in range(args.steps):
    t0 = time.perf_counter()
    dg.neighbors_batch(dmgraph.BOTH, 2, 1, q_off, q_src, out=bufs)
    ts.append(time.perf_counter() - t0)
out["sharded_batch"] = {"queries": args.queries, "results": int(ids.size), "checked_against": "oracle", "matches": True,
                        "e2e_ms": 1e3 * float(np.median(ts)), "queries_per_s": args.queries / float(np.median(ts)),
                        "note": "sizing + fetch traversals, pageable host buffers, one host thread per device"}
dg.release()

# (2) partitioned closure with NCCL allgather per level -- every level word against the oracle
t0 = time.perf_counter()
big = synth.big_graph_fast(args.vertices, args.edges, cap=2000, entry_frac=0.01, seed=synth.SEED_C5)
gen_s = time.perf_counter() - t0
t0 = time.perf_counter()
pg = dmgraph.DistGraph(dist, big.n_vertices, big.src, big.dst, big.entry, partitioned=True)
upload_s = time.perf_counter() - t0
o = ora.CsrOracle(big.n_vertices, big.src, big.dst, big.entry)
wlvl, wreached, wte = o.closure(0, big.entry)
out["partitioned_closure"] = {"vertices": big.n_vertices, "edges_drawn": big.n_edges, "reached": int(wreached),
                              "checked_against": "oracle", "generate_s": gen_s, "upload_s": upload_s,
                              "definitional_edges": int(wte)}
for name, env in (("peer_memory", None), ("nccl_allgather", "1")):
    if env:
        os.environ["DMG_DIST_NCCL"] = env
    else:
        os.environ.pop("DMG_DIST_NCCL", None)
    lvl, reached, st = pg.closure(dmgraph.OUT, big.entry)
    assert reached == wreached and np.array_equal(lvl, wlvl), name
    best = None
    for _ in range(args.steps):
        _l, _r, s = pg.closure(dmgraph.OUT, big.entry)
        assert _r == wreached
        if best is None or s["device_ms"] < best["device_ms"]:
            best = s
    assert np.array_equal(_l, wlvl), name + " (repeat)"
    out["partitioned_closure"][name] = {"matches": True, **best,
                                        "gteps_definitional": wte / (best["device_ms"] * 1e-3) / 1e9}
os.environ.pop("DMG_DIST_NCCL", None)
pg.release()

# (3) a sparse graph (average degree 3) from few sources, both directions: many push levels, vertex counts that
# do not divide by 32 * N, sources in one rank's range only
rng = np.random.Generator(np.random.PCG64(77))
Vs = 1_000_003
ss = rng.integers(0, Vs, 3 * Vs).astype(np.uint32)
dd = rng.integers(0, Vs, 3 * Vs).astype(np.uint32)
srcs = rng.integers(0, 5000, 7).astype(np.uint32)
sp = dmgraph.DistGraph(dist, Vs, ss, dd, srcs, partitioned=True)
o = ora.CsrOracle(Vs, ss, dd, srcs)
out["sparse_closure"] = {"vertices": Vs, "edges_drawn": int(ss.size), "checked_against": "oracle"}
for direction, dname in ((dmgraph.OUT, "out"), (dmgraph.IN, "in")):
    wlvl, wreached, _ = o.closure(direction, srcs)
    for name, env in (("peer_memory", None), ("nccl_allgather", "1")):
        if env:
            os.environ["DMG_DIST_NCCL"] = env
        else:
            os.environ.pop("DMG_DIST_NCCL", None)
        for _ in range(2):
            lvl, reached, st = sp.closure(direction, srcs)
            assert reached == wreached and np.array_equal(lvl, wlvl), (dname, name)
        out["sparse_closure"][dname + "_" + name] = {"matches": True, "reached": int(reached), **st}
os.environ.pop("DMG_DIST_NCCL", None)
# (4) exact analysisOrder of the partitioned sparse graph (7 entry points) against the oracle's FIFO order
t0 = time.perf_counter()
got = sp.analysis_order()
t_order = time.perf_counter() - t0
assert np.array_equal(got, o.analysis_order())
out["partitioned_analysis_order"] = {"vertices": Vs, "checked_against": "oracle", "matches": True, "wall_ms": 1e3 * t_order}
sp.release()
dist.close()
print(json.dumps(out), flush=True)
