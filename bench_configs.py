#!/usr/bin/env python3
"""Measurements for the BASELINE configs that bench.py (config 2, the headline) does not cover.
One JSON line per run; results are kept under profiles/.

  python bench_configs.py c3 [--names N] [--queries Q]     search_project scan (S1)
  python bench_configs.py c4 [--vertices V] [--edges E]    adjacency rows + directed k-hop on one big graph
  python bench_configs.py c4mix [--ops N]                  the SURVEY 8d config-4 DSL mix in one dmg_query_batch
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


def run_c4mix(args):
    """BASELINE config 4 as SURVEY.md 8d states it: one 1 M / 30 M graph, a batch of resolved DSL
    queries in the stated mix (40 % dependencies, 20 % dependents, 15 % overview, 10 % k-hop OUT
    k in {2,3,4}, 5 % simple-name resolution, 5 % ?check, 3 % entrypoints, 2 % endpoints) through ONE
    dmg_query_batch call with host buffers."""
    from oracle import oracle as ora
    eng = dmgraph.Engine(0)
    t0 = time.perf_counter()
    sg = synth.big_graph(args.vertices, args.edges, cap=500, entry_frac=0.01, seed=synth.SEED_C4, with_names=True)
    V = sg.n_vertices
    off, arena = synth.pack_names(sg.names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    # 10 methods per vertex, 8 lower-case bytes each: 4-letter verb + 4 hex digits of (7 v + j)
    verbs = np.frombuffer(b"findsavelistloadsendreadpushpullopenstop", dtype=np.uint8).reshape(10, 4)
    num = ((np.arange(V, dtype=np.int64)[:, None] * 7 + np.arange(10)[None, :]) & 0xFFFF)
    hexd = np.frombuffer(b"0123456789abcdef", dtype=np.uint8)
    digits = hexd[(num[:, :, None] >> np.array([12, 8, 4, 0])[None, None, :]) & 15]
    mbytes = np.concatenate([np.broadcast_to(verbs[None, :, :], (V, 10, 4)), digits], axis=2).reshape(-1)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    g = dmgraph.Graph(eng, V, sg.src, sg.dst, sg.entry, name_off=off, name_bytes=arena, lower_bytes=low)
    g.set_methods_packed(np.arange(V + 1, dtype=np.uint32) * 10, np.arange(10 * V + 1, dtype=np.uint32) * 8, mbytes)
    rng = np.random.Generator(np.random.PCG64(synth.SEED_C4 + 2))
    ctrl = rng.random(V) < 0.01                                  # 1 % controllers: all their methods are endpoints
    g.set_method_flags(np.repeat(ctrl.astype(np.uint8), 10))
    upload_s = time.perf_counter() - t0

    def make_ops(n):
        r = np.random.Generator(np.random.PCG64(synth.SEED_C4 + 3))
        u = r.random(n)
        edges_ = np.cumsum([0.40, 0.20, 0.15, 0.10, 0.05, 0.05, 0.03, 0.02])
        kind = np.searchsorted(edges_, u, side="right").clip(0, 7)
        opcode = np.array([dmgraph.OP_DEPENDENCIES, dmgraph.OP_DEPENDENTS, dmgraph.OP_OVERVIEW, dmgraph.OP_KHOP,
                           dmgraph.OP_RESOLVE_NAME, dmgraph.OP_METHOD, dmgraph.OP_ENTRYPOINTS, dmgraph.OP_ENDPOINTS],
                          np.uint32)[kind]
        ops = np.zeros(n, dmgraph.Graph.QUERY_OP_DTYPE)
        ops["opcode"] = opcode
        ops["vertex"] = r.integers(0, V, n)
        ops["mode"] = dmgraph.OUT
        ops["k"] = np.where(opcode == dmgraph.OP_KHOP, r.integers(2, 5, n), 1)
        text = bytearray()
        for i in np.flatnonzero((opcode == dmgraph.OP_RESOLVE_NAME) | (opcode == dmgraph.OP_METHOD)):
            v = int(ops["vertex"][i])
            if opcode[i] == dmgraph.OP_RESOLVE_NAME:
                t = sg.names[v].rsplit(".", 1)[-1].lower().encode()
            elif r.random() < 0.8:
                j = int(r.integers(0, 10))
                t = bytes(mbytes[(v * 10 + j) * 8:(v * 10 + j + 1) * 8])
            else:
                t = b"nosuchmethod"
            ops["text_begin"][i] = len(text)
            text += t
            ops["text_end"][i] = len(text)
        return ops, np.frombuffer(bytes(text) + b"\0", dtype=np.uint8).copy(), kind

    n = args.ops
    while True:
        ops, text, kind = make_ops(n)
        t0 = time.perf_counter()
        need = g.query_batch_packed(ops, text)
        size_s = time.perf_counter() - t0
        if need <= args.max_results or n <= 1024:
            break
        n //= 2
    off_o, _p0 = eng.pinned(np.uint64, n + 1)
    ids, _p1 = eng.pinned(np.uint32, max(need, 1))
    aux, _p2 = eng.pinned(np.uint8, max(need, 1))
    eng.profile(True)
    ts = []
    for it in range(1 + args.steps):
        eng.profile_reset()
        t0 = time.perf_counter()
        got = g.query_batch_packed(ops, text, out=(off_o, ids, aux))
        ts.append(time.perf_counter() - t0)
    assert got == need
    kt = kernels(eng)
    kms = sum(v["ms_total"] for v in kt.values())
    # spot check against the csr restatement
    o = ora.CsrOracle(V, sg.src, sg.dst, sg.entry)
    checked = 0
    for i in np.random.Generator(np.random.PCG64(1)).choice(n, min(n, 300), replace=False):
        a, b = int(off_o[i]), int(off_o[i + 1])
        v, op = int(ops["vertex"][i]), int(ops["opcode"][i])
        if op == dmgraph.OP_DEPENDENCIES:
            assert ids[a:b].tolist() == o.adjacency_batch(0, [v])[1].tolist()
        elif op == dmgraph.OP_DEPENDENTS:
            assert ids[a:b].tolist() == o.adjacency_batch(1, [v])[1].tolist()
        elif op == dmgraph.OP_KHOP:
            _o, wi, wh, _te = o.khop_batch(0, int(ops["k"][i]), 0, [0, 1], [v])
            assert ids[a:b].tolist() == wi.tolist() and aux[a:b].tolist() == wh.tolist()
        elif op == dmgraph.OP_ENTRYPOINTS:
            assert ids[a:b].tolist() == o.entry_points().tolist()
        else:
            continue
        checked += 1
    names_k = ["dependencies", "dependents", "overview", "khop", "resolve_name", "method", "entrypoints", "endpoints"]
    line = {"config": "c4mix", "workload": "one graph %d V / %d E, %d DSL ops in the SURVEY 8d mix, one dmg_query_batch call "
                                           "(pinned host buffers)" % (V, g.info["n_edges"], n),
            "ops_requested": args.ops, "ops_run": n, "results": int(need),
            "op_counts": {names_k[k]: int((kind == k).sum()) for k in range(8)},
            "metric": "dsl_ops_per_sec", "unit": "ops/s",
            "e2e": {"value": n / float(np.median(ts[1:])), "ms_per_batch": 1e3 * float(np.median(ts[1:])),
                    "sizing_call_ms": 1e3 * size_s, "result_bytes": int(5 * need + 8 * (n + 1))},
            "kernel_ms_per_batch": kms, "kernels": kt, "rows_checked_against_oracle": checked,
            "note": "k-hop groups run as device batches (count pass + fetch pass), rows are assembled on the device; "
                    "e2e minus kernel time is host grouping, per-call allocations and the PCIe copy of the result",
            "generate_s": gen_s, "upload_s": upload_s}
    print(json.dumps(line), flush=True)
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
    ap.add_argument("config", choices=["c3", "c4", "c4mix", "c5"])
    ap.add_argument("--names", type=int, default=10_000_000)
    ap.add_argument("--queries", type=int, default=16384)
    ap.add_argument("--vertices", type=int, default=1_000_000)
    ap.add_argument("--edges", type=int, default=30_000_000)
    ap.add_argument("--ops", type=int, default=262_144)
    ap.add_argument("--max-results", type=int, default=600_000_000,
                    help="c4mix: halve the batch until the packed result has at most this many entries")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    args = ap.parse_args()
    {"c3": run_c3, "c4": run_c4, "c4mix": run_c4mix, "c5": run_c5}[args.config](args)


if __name__ == "__main__":
    main()
