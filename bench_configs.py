#!/usr/bin/env python3
"""Measurements for the BASELINE configs that bench.py (config 2, the headline) does not cover.
One JSON line per run; results are kept under profiles/.

  python bench_configs.py c3 [--names N] [--queries Q]     search_project scan (S1)
  python bench_configs.py c4 [--vertices V] [--edges E]    adjacency rows + directed k-hop on one big graph
  python bench_configs.py c5 [--vertices V] [--edges E]    closure levels + exact analysisOrder

Per-kernel times are CUDA events on the library's launch stream (dmg_profile_*); e2e is the
wall clock of the host-pointer C-ABI call.  The CPU numbers are the oracle's restatements on the
box's host cores (test infrastructure used as the timed baseline only).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "domain-mcp-server_b200"))
sys.path.insert(0, ROOT)

import dmgraph  # noqa: E402
from dmgraph import synth  # noqa: E402

HBM = 6540.8
try:
    HBM = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def kernels(eng):
    return {k: {"launches": int(v[0]), "ms_total": v[1]} for k, v in eng.profile_table().items()}


def run_c3(args):
    from oracle import oracle as ora
    eng = dmgraph.Engine(0)
    t0 = time.perf_counter()
    off, arena = synth.catalog_names(args.names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    queries = synth.search_queries(off, arena, args.queries)
    gen_s = time.perf_counter() - t0
    g = dmgraph.Graph(eng, off.size - 1, [], [], [], name_off=off, name_bytes=arena, lower_bytes=low)
    qbytes = sum(len(q) for q in queries)
    # sizing call, then timed calls with an exact capacity
    o, ids = g.search_batch(dmgraph.SEARCH_CONTAINS_ALL, queries)
    H = int(o[-1])
    times, dev = [], []
    eng.profile(True)
    eng.profile_reset()
    for it in range(args.warmup + args.steps):
        eng.flush_l2()
        if it == args.warmup:
            eng.profile_reset()
        t0 = time.perf_counter()
        o2, ids2 = g.search_batch(dmgraph.SEARCH_CONTAINS_ALL, queries, cap=H)
        dt = time.perf_counter() - t0
        if it >= args.warmup:
            times.append(dt)
            dev.append(eng.last_stats()["device_ms"])
    kt = kernels(eng)
    assert np.array_equal(o, o2) and np.array_equal(ids, ids2)
    kernel_ms = sum(v["ms_total"] for v in kt.values()) / args.steps
    A = int(off[-1])
    N = off.size - 1
    alg = A + 4 * (N + 1) + qbytes + 4 * H + 4 * (len(queries) + 1)      # SURVEY.md 8d, single-pass minimum
    scan_ms = kt.get("search_scan", {"ms_total": 0})["ms_total"] / args.steps
    emit_ms = kt.get("search_emit", {"ms_total": 0})["ms_total"] / args.steps
    # CPU: csr-style (memmem over the lowered arena) on a query sample, all threads
    threads = ora.hardware_threads()
    ns = max(threads, min(len(queries), 4 * threads))
    t0 = time.perf_counter()
    ora.search_batch(off, low, queries[:ns], 0, threads=threads)
    cpu_s = time.perf_counter() - t0
    line = {"config": "c3", "workload": "search_project scan: %d names (%.0f MB arena), %d queries, %d hits" % (N, A / 1e6, len(queries), H),
            "metric": "search_queries_per_sec", "unit": "queries/s",
            "value_kernels": len(queries) / (kernel_ms * 1e-3), "kernel_ms_per_batch": kernel_ms,
            "e2e": {"value": len(queries) / np.median(times), "ms_per_batch": 1e3 * float(np.median(times)),
                    "device_ms": float(np.median(dev)), "d2h_bytes": 4 * H + 8 * (len(queries) + 1)},
            "roofline": {"bound": "hbm", "unit": "GB/s", "peak": HBM, "algorithmic_bytes": alg,
                         "achieved_whole_batch": alg / (kernel_ms * 1e-3) / 1e9,
                         "frac_whole_batch": alg / (kernel_ms * 1e-3) / 1e9 / HBM,
                         "match_pass_GBs": (A + 4 * N) / (scan_ms * 1e-3) / 1e9 if scan_ms else None,
                         "emit_pass_GBs": (A + 4 * N + 4 * H) / (emit_ms * 1e-3) / 1e9 if emit_ms else None,
                         "effective_scan_rate_TBs": len(queries) * A / (kernel_ms * 1e-3) / 1e12},
            "kernels": kt,
            "cpu_baseline": {"kind": "port", "cores": threads, "unit": "queries/s", "value": ns / cpu_s,
                             "sample": "%d of %d queries over all %d names, memmem over the pre-lowered arena (csr "
                                       "restatement; the string-allocating ref variant is slower)" % (ns, len(queries), N)},
            "generate_s": gen_s}
    print(json.dumps(line), flush=True)
    g.release()
    eng.close()


def run_c4(args):
    from oracle import oracle as ora
    eng = dmgraph.Engine(0)
    t0 = time.perf_counter()
    sg = synth.big_graph(args.vertices, args.edges, cap=500, entry_frac=0.01, seed=synth.SEED_C4)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    g = dmgraph.Graph(eng, sg.n_vertices, sg.src, sg.dst, sg.entry)
    upload_s = time.perf_counter() - t0
    rng = np.random.Generator(np.random.PCG64(synth.SEED_C4 + 1))
    nq = args.queries
    ids = rng.integers(0, sg.n_vertices, nq).astype(np.uint32)
    out = {"config": "c4", "workload": "one graph %d V / %d E, %d ids per primitive batch" % (sg.n_vertices, g.info["n_edges"], nq),
           "generate_s": gen_s, "upload_s": upload_s, "primitives": {}}
    eng.profile(True)

    def timed(name, fn, bytes_fn):
        fn()
        eng.profile_reset()
        ts = []
        for _ in range(args.steps):
            eng.flush_l2()
            t0 = time.perf_counter()
            r = fn()
            ts.append(time.perf_counter() - t0)
        st = eng.last_stats()
        kt = kernels(eng)
        kms = sum(v["ms_total"] for v in kt.values()) / args.steps
        b = bytes_fn(st)
        out["primitives"][name] = {"queries_per_s_kernels": nq / (kms * 1e-3), "kernel_ms": kms,
                                   "e2e_ms": 1e3 * float(np.median(ts)), "results": int(st["results"]),
                                   "edges_traversed": int(st["edges_traversed"]),
                                   "gteps": st["edges_traversed"] / (kms * 1e-3) / 1e9,
                                   "algorithmic_bytes": int(b), "achieved_GBs": b / (kms * 1e-3) / 1e9,
                                   "frac_of_hbm": b / (kms * 1e-3) / 1e9 / HBM, "kernels": kt}
        return r

    q_off = np.arange(nq + 1, dtype=np.uint32)
    timed("dependencies (adjacency OUT)", lambda: g.adjacency_batch(dmgraph.OUT, ids),
          lambda st: 12 * nq + 8 * st["results"] + 8 * nq)
    timed("dependents (adjacency IN)", lambda: g.adjacency_batch(dmgraph.IN, ids),
          lambda st: 12 * nq + 8 * st["results"] + 8 * nq)
    for k in (1, 2, 3):
        timed("k-hop OUT k=%d" % k, lambda k=k: g.neighbors_batch(dmgraph.OUT, k, 0, q_off[:nq // 8 + 1], ids[:nq // 8]),
              lambda st: 8 * st["rows_expanded"] + 4 * st["edges_traversed"] + 4 * (nq // 8) + 5 * st["results"])
    print(json.dumps(out), flush=True)
    g.release()
    eng.close()


def run_c5(args):
    from oracle import oracle as ora
    eng = dmgraph.Engine(0)
    t0 = time.perf_counter()
    sg = synth.big_graph(args.vertices, args.edges, cap=2000, entry_frac=0.01, seed=synth.SEED_C5)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    g = dmgraph.Graph(eng, sg.n_vertices, sg.src, sg.dst, sg.entry)
    upload_s = time.perf_counter() - t0
    eng.profile(True)
    out = {"config": "c5", "workload": "closure on %d V / %d E (single GPU, replicated)" % (sg.n_vertices, g.info["n_edges"]),
           "generate_s": gen_s, "upload_s": upload_s}
    V = sg.n_vertices
    for name, fn in [("closure_from_entry_points", lambda: g.closure(dmgraph.OUT, sg.entry)),
                     ("analysis_order_exact", lambda: g.analysis_order())]:
        fn()
        eng.profile_reset()
        ts = []
        for _ in range(args.steps):
            eng.flush_l2()
            t0 = time.perf_counter()
            r = fn()
            ts.append(time.perf_counter() - t0)
        st = eng.last_stats()
        kt = kernels(eng)
        kms = sum(v["ms_total"] for v in kt.values()) / args.steps
        te = st["edges_traversed"]
        reached = st["results"]
        b = 8 * st["rows_expanded"] + 4 * te + 4 * V
        out[name] = {"kernel_ms": kms, "e2e_ms": 1e3 * float(np.median(ts)), "edges_traversed": int(te),
                     "gteps": te / (kms * 1e-3) / 1e9 if te else None, "reached": int(reached),
                     "algorithmic_bytes": int(b), "achieved_GBs": b / (kms * 1e-3) / 1e9,
                     "frac_of_hbm": b / (kms * 1e-3) / 1e9 / HBM, "kernels": kt}
    # CPU csr closure for context
    o = ora.CsrOracle(V, sg.src, sg.dst, sg.entry)
    t0 = time.perf_counter()
    lvl, reached, te = o.closure(0, sg.entry)
    out["cpu_csr_closure"] = {"seconds": time.perf_counter() - t0, "cores": 1, "reached": int(reached), "edges": int(te)}
    glvl, greached = g.closure(dmgraph.OUT, sg.entry)
    assert greached == reached and np.array_equal(glvl, lvl)
    print(json.dumps(out), flush=True)
    g.release()
    eng.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["c3", "c4", "c5"])
    ap.add_argument("--names", type=int, default=10_000_000)
    ap.add_argument("--queries", type=int, default=16384)
    ap.add_argument("--vertices", type=int, default=1_000_000)
    ap.add_argument("--edges", type=int, default=30_000_000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    args = ap.parse_args()
    {"c3": run_c3, "c4": run_c4, "c5": run_c5}[args.config](args)


if __name__ == "__main__":
    main()
