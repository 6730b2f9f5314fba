#!/usr/bin/env python3
"""Measurements for the BASELINE configs beside the headline (config 2): the HBM-bound ones.
bench.py calls these at N = 1 and puts their dicts under "configs" in its JSON line; run alone,
one JSON line per config is printed (kept under profiles/).

  python bench_configs.py c3 [--names N] [--queries Q]     search_project scan (S1)
  python bench_configs.py c4 [--vertices V] [--edges E]    adjacency rows, directed k-hop, the DSL mix
  python bench_configs.py c5 [--vertices V] [--edges E]    closure levels + exact analysisOrder

Every block reports: kernel time per batch (CUDA events on the library's launch stream, L2 flushed
before each timed call), SURVEY.md 8d algorithmic bytes and their fraction of the measured HBM peak,
e2e through the host-pointer C-ABI call (pinned host buffers, copies inside the timed region), an
oracle check of the very results that were timed (`parity_checked` = rows compared), and a CPU figure
from the oracle's restatements on the box's host cores (test infrastructure used as checker and
timed baseline only).
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


def kernels(eng, steps=1):
    return {k: {"launches": int(v[0]), "ms_per_batch": v[1] / steps} for k, v in eng.profile_table().items()}


def roofline(alg_bytes, kernel_ms, **extra):
    ach = alg_bytes / (kernel_ms * 1e-3) / 1e9 if kernel_ms else None
    d = {"bound": "hbm", "unit": "GB/s", "peak": HBM, "algorithmic_bytes": int(alg_bytes), "kernel_ms": kernel_ms,
         "achieved": ach, "frac": ach / HBM if ach else None}
    d.update(extra)
    return d


def timed_calls(eng, fn, steps, warmup=1):
    """fn() is one host-pointer C-ABI call.  Returns (kernel ms per call from the per-launch events,
    per-kernel table, median wall ms, stats of the last call, last return value)."""
    r = None
    for _ in range(warmup):
        r = fn()
    eng.profile(True)
    eng.profile_reset()
    wall = []
    for _ in range(steps):
        eng.flush_l2()
        t0 = time.perf_counter()
        r = fn()
        wall.append(1e3 * (time.perf_counter() - t0))
    st = eng.last_stats()
    kt = kernels(eng, steps)
    eng.profile(False)
    return sum(v["ms_per_batch"] for v in kt.values()), kt, float(np.median(wall)), st, r


# ---------------------------------------------------------------------------------------------------
def run_c3(eng, names=10_000_000, n_queries=16384, steps=3, check_queries=4096, cpu_budget_s=10.0):
    from oracle import oracle as ora
    t0 = time.perf_counter()
    off, arena = synth.catalog_names(names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    queries = synth.search_queries(off, arena, n_queries)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    g = dmgraph.Graph(eng, off.size - 1, [], [], [], name_off=off, name_bytes=arena, lower_bytes=low)
    upload_s = time.perf_counter() - t0
    q_off = np.zeros(len(queries) + 1, np.uint32)
    np.cumsum([len(q) for q in queries], out=q_off[1:])
    qb = np.frombuffer(b"".join(queries) + b"\0", dtype=np.uint8).copy()
    # sizing call (capacity 0 answers DMG_ECAP with `needed`), then timed calls into pinned buffers
    import ctypes as C
    need = C.c_uint64(0)
    off0 = np.zeros(len(queries) + 1, np.uint64)
    g.L.dmg_search_batch(g.h, dmgraph.SEARCH_CONTAINS_ALL, 0, g.V, len(queries), q_off.ctypes.data, qb.ctypes.data,
                         off0.ctypes.data, None, 0, C.byref(need))
    H = int(need.value)
    h_off, p1 = eng.pinned(np.uint64, len(queries) + 1)
    h_ids, p2 = eng.pinned(np.uint32, max(H, 1))
    kms, kt, wall_ms, st, _r = timed_calls(
        eng, lambda: g.search_batch(dmgraph.SEARCH_CONTAINS_ALL, (q_off, qb), out=(h_off, h_ids)), steps)
    A, N, Q = int(off[-1]), off.size - 1, len(queries)
    alg = A + 4 * (N + 1) + int(q_off[-1]) + 4 * H + 4 * (Q + 1)          # SURVEY.md 8d, single-pass minimum
    # oracle check of the timed result: Aho-Corasick over ALL names for a seeded sample of the queries
    threads = ora.hardware_threads()
    rng = np.random.Generator(np.random.PCG64(33))
    sample = np.sort(rng.choice(Q, min(Q, check_queries), replace=False))
    t0 = time.perf_counter()
    wo, wids = ora.search_batch_ac(off, low, [queries[i] for i in sample], threads=threads)
    check_s = time.perf_counter() - t0
    for j, i in enumerate(sample):
        a, b = int(h_off[i]), int(h_off[i + 1])
        assert b - a == int(wo[j + 1] - wo[j]) and np.array_equal(h_ids[a:b], wids[int(wo[j]):int(wo[j + 1])]), \
            "c3: hits of query %d differ from the oracle" % i
    # CPU baseline: the restatement's loop (one contains() per query and name) on a bounded sample
    ns = threads
    t0 = time.perf_counter()
    ora.search_batch(off, low, queries[:ns], 0, threads=threads)
    dt = time.perf_counter() - t0
    ns2 = int(min(Q, max(threads, (cpu_budget_s / max(dt, 1e-6)) * ns)) // threads * threads)
    if ns2 > ns:
        t0 = time.perf_counter()
        ora.search_batch(off, low, queries[:ns2], 0, threads=threads)
        dt, ns = time.perf_counter() - t0, ns2
    out = {"workload": "c3: search_project scan, %d names (%.0f MB arena), %d queries, %d ordered hits" % (N, A / 1e6, Q, H),
           "metric": "search_queries_per_sec", "unit": "queries/s", "value": Q / (kms * 1e-3),
           "kernel_ms_per_batch": kms, "kernels": kt,
           "roofline": roofline(alg, kms, effective_scan_rate_TBs=Q * A / (kms * 1e-3) / 1e12,
                                note="bytes = arena + offsets + query text + 4 B per hit + row offsets (one pass)"),
           "e2e": {"value": Q / (wall_ms * 1e-3), "unit": "queries/s", "ms_per_batch": wall_ms,
                   "device_ms": st["device_ms"], "h2d_bytes": int(st["h2d_bytes"]), "d2h_bytes": int(st["d2h_bytes"]),
                   "host_buffers": "pinned"},
           "parity_checked": int(sample.size), "parity_against": "oracle (Aho-Corasick form, all names)",
           "parity_check_s": check_s,
           "cpu_baseline": {"kind": "port", "cores": threads, "unit": "queries/s", "value": ns / dt,
                            "sample": "%d of %d queries over all %d names: contains() per (query, name) on the "
                                      "pre-lowered arena, %.1f s" % (ns, Q, N, dt)},
           "generate_s": gen_s, "upload_s": upload_s}
    eng.pinned_free(p1)
    eng.pinned_free(p2)
    g.release()
    return out


# ---------------------------------------------------------------------------------------------------
def run_c4(eng, vertices=1_000_000, edges=30_000_000, n_ids=262_144, n_khop=32_768, n_ops=262_144, steps=3,
           ops_per_call=16_384, check_rows=4096):
    from oracle import checks
    from oracle import oracle as ora
    threads = ora.hardware_threads()
    t0 = time.perf_counter()
    sg = synth.big_graph(vertices, edges, cap=500, entry_frac=0.01, seed=synth.SEED_C4, with_names=True)
    V = sg.n_vertices
    off, arena = synth.pack_names(sg.names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    mo, no, mbytes = synth.method_catalog(V)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    g = dmgraph.Graph(eng, V, sg.src, sg.dst, sg.entry, name_off=off, name_bytes=arena, lower_bytes=low)
    g.set_methods_packed(mo, no, mbytes)
    rng = np.random.Generator(np.random.PCG64(synth.SEED_C4 + 2))
    flags = np.repeat((rng.random(V) < 0.01).astype(np.uint8), 10)      # 1 % controllers: their methods are endpoints
    g.set_method_flags(flags)
    endpoint_methods = np.flatnonzero(flags).astype(np.uint32)
    upload_s = time.perf_counter() - t0
    o = ora.CsrOracle(V, sg.src, sg.dst, sg.entry)
    rng = np.random.Generator(np.random.PCG64(synth.SEED_C4 + 1))
    ids = rng.integers(0, V, n_ids).astype(np.uint32)
    out = {"workload": "c4: one graph %d V / %d E; %d ids per adjacency batch, %d queries per k-hop batch, "
                       "%d DSL ops in the SURVEY 8d mix" % (V, g.info["n_edges"], n_ids, n_khop, n_ops),
           "generate_s": gen_s, "upload_s": upload_s, "primitives": {}}
    checked = 0

    def record(name, nq, kms, kt, wall_ms, st, alg):
        out["primitives"][name] = {
            "queries_per_s": nq / (kms * 1e-3), "kernel_ms": kms, "kernels": kt, "results": int(st["results"]),
            "edges_traversed": int(st["edges_traversed"]), "gteps": st["edges_traversed"] / (kms * 1e-3) / 1e9,
            "roofline": roofline(alg, kms),
            "e2e": {"ms": wall_ms, "h2d_bytes": int(st["h2d_bytes"]), "d2h_bytes": int(st["d2h_bytes"])}}

    pins = []

    def pinned_out(n_rows, entries, with_hop):
        bufs = [eng.pinned(np.uint64, n_rows + 1), eng.pinned(np.uint32, max(entries, 1))]
        if with_hop:
            bufs.append(eng.pinned(np.uint8, max(entries, 1)))
        pins.extend(p for _a, p in bufs)
        return tuple(a for a, _p in bufs)

    for name, d in (("dependencies", dmgraph.OUT), ("dependents", dmgraph.IN)):
        n_res = int(g.adjacency_batch(d, ids)[0][-1])                       # sizing, untimed
        bufs = pinned_out(n_ids, n_res, False)
        kms, kt, wall, st, r = timed_calls(eng, lambda d=d, bufs=bufs: g.adjacency_batch(d, ids, out=bufs), steps)
        # SURVEY 8d per id: 8 B row pointers + 4 B per entry read, 4 B per entry + 8 B offset written, 4 B id
        record(name, n_ids, kms, kt, wall, st, 20 * n_ids + 8 * st["results"])
        woff, wids = o.adjacency_batch(d, ids)
        assert np.array_equal(r[0], woff) and np.array_equal(r[1], wids), "c4 %s differs from the oracle" % name
        checked += n_ids
    q_off = np.arange(n_khop + 1, dtype=np.uint32)
    ksrc = ids[:n_khop]
    for k in (1, 2, 3):
        n_res = int(g.neighbors_batch(dmgraph.OUT, k, 0, q_off, ksrc)[0][-1])   # sizing, untimed
        bufs = pinned_out(n_khop, n_res, True)
        kms, kt, wall, st, r = timed_calls(
            eng, lambda k=k, bufs=bufs: g.neighbors_batch(dmgraph.OUT, k, 0, q_off, ksrc, out=bufs), steps)
        record("khop_out_k%d" % k, n_khop, kms, kt, wall, st,
               8 * st["rows_expanded"] + 4 * st["edges_traversed"] + 4 * n_khop + 5 * st["results"])
        rows = np.sort(np.random.Generator(np.random.PCG64(k)).choice(n_khop, min(n_khop, 1024), replace=False))
        checked += checks.check_khop_rows(o, 0, k, 0, q_off, ksrc, r[0], r[1], r[2], rows=rows, threads=threads)

    # the DSL mix at the stated batch size, in calls of ops_per_call operations (one call's packed result
    # has to fit the host's buffers: the endpoint / entry-point rows alone are ~100 k entries each)
    ops, text, kind = synth.c4_mix_ops(n_ops, sg.names, mbytes, dmgraph.Graph.QUERY_OP_DTYPE)
    chunks = [(a, min(n_ops, a + ops_per_call)) for a in range(0, n_ops, ops_per_call)]
    t0 = time.perf_counter()
    needs = [g.query_batch_packed(ops[a:b], text) for a, b in chunks]
    sizing_s = time.perf_counter() - t0
    cap = max(max(needs), 1)
    h_off, p1 = eng.pinned(np.uint64, ops_per_call + 1)
    h_ids, p2 = eng.pinned(np.uint32, cap)
    h_aux, p3 = eng.pinned(np.uint8, cap)
    rng = np.random.Generator(np.random.PCG64(5))
    want_rows = np.sort(rng.choice(n_ops, min(n_ops, check_rows), replace=False))
    # one untimed call first (as every other timed region here has its warm-up): the first fetch grows the
    # context's buffers to this workload's sizes
    a0, b0 = chunks[0]
    g.query_batch_packed(ops[a0:b0], text, out=(h_off[:b0 - a0 + 1], h_ids, h_aux))
    eng.profile(True)
    eng.profile_reset()
    mix_checked, d2h, h2d = 0, 0, 0
    t0 = time.perf_counter()
    call_s = 0.0
    call_ms = []
    for (a, b), need in zip(chunks, needs):
        t1 = time.perf_counter()
        got = g.query_batch_packed(ops[a:b], text, out=(h_off[:b - a + 1], h_ids, h_aux))
        call_ms.append(1e3 * (time.perf_counter() - t1))
        call_s += 1e-3 * call_ms[-1]
        assert got == need
        st = eng.last_stats()
        d2h += st["d2h_bytes"]
        h2d += st["h2d_bytes"]
        rows = want_rows[(want_rows >= a) & (want_rows < b)] - a
        mix_checked += checks.check_query_mix(o, ops[a:b], text, h_off, h_ids, h_aux, off, low, mbytes,
                                              endpoint_methods, rows=rows, threads=threads)
    kt = kernels(eng)
    eng.profile(False)
    kms = sum(v["ms_per_batch"] for v in kt.values())
    # the same 16 calls from TWO host threads (the engine hands every call its own stream and workspace, as the
    # MCP server's concurrent requests would use it): one call's kernels run under the other's copy to the host
    import threading
    bufs2 = [(h_off, h_ids, h_aux)]
    q_off2, q1 = eng.pinned(np.uint64, ops_per_call + 1)
    q_ids2, q2 = eng.pinned(np.uint32, cap)
    q_aux2, q3 = eng.pinned(np.uint8, cap)
    bufs2.append((q_off2, q_ids2, q_aux2))
    errs = []

    def worker(w):
        try:
            o_, i_, a_ = bufs2[w]
            for ci in range(w, len(chunks), 2):
                a, b = chunks[ci]
                assert g.query_batch_packed(ops[a:b], text, out=(o_[:b - a + 1], i_, a_)) == needs[ci]
        except Exception as ex:      # surfaces in the block's JSON instead of killing the thread silently
            errs.append("%s: %s" % (type(ex).__name__, ex))

    t1 = time.perf_counter()
    th = [threading.Thread(target=worker, args=(w,)) for w in range(2)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    two_s = time.perf_counter() - t1
    for q in (q1, q2, q3):
        eng.pinned_free(q)
    total = int(sum(needs))
    names_k = [k for k, _ in synth.C4_MIX]
    out["mix"] = {"ops": n_ops, "calls": len(chunks), "results": total,
                  "op_counts": {names_k[k]: int((kind == k).sum()) for k in range(8)},
                  "metric": "dsl_ops_per_sec", "unit": "ops/s", "value": n_ops / (kms * 1e-3), "kernel_ms": kms,
                  "kernels": kt,
                  "roofline": roofline(5 * total + 8 * (n_ops + len(chunks)) + 32 * n_ops, kms,
                                       note="bytes = packed result (4 B id + 1 B aux per entry) + offsets + ops; "
                                            "the rows are copies of resident CSR rows / k-hop rows"),
                  "e2e": {"value": n_ops / call_s, "unit": "ops/s", "ms": 1e3 * call_s, "h2d_bytes": int(h2d),
                          "d2h_bytes": int(d2h), "host_buffers": "pinned", "sizing_calls_s": sizing_s,
                          "ms_per_call": [round(x, 2) for x in call_ms], "warmup_calls": 1,
                          "two_host_threads": {"ms": 1e3 * two_s, "value": n_ops / two_s, "unit": "ops/s",
                                               "errors": errs,
                                               "note": "the same calls issued from two threads with their own pinned "
                                                       "buffers: kernels of one call overlap the other's D2H copy"}},
                  "parity_checked": int(mix_checked), "opcodes_checked": sorted({names_k[k] for k in kind[want_rows]})}
    out["parity_checked"] = int(checked + mix_checked)
    out["parity_against"] = "oracle csr (adjacency: whole batches; k-hop and mix: seeded row samples, every opcode)"
    for p in [p1, p2, p3] + pins:
        eng.pinned_free(p)
    g.release()
    return out


# ---------------------------------------------------------------------------------------------------
def run_c5(eng, vertices=4_000_000, edges=100_000_000, steps=3):
    from oracle import oracle as ora
    t0 = time.perf_counter()
    sg = synth.big_graph_fast(vertices, edges, cap=2000, entry_frac=0.01, seed=synth.SEED_C5)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    g = dmgraph.Graph(eng, sg.n_vertices, sg.src, sg.dst, sg.entry)
    upload_s = time.perf_counter() - t0
    V = sg.n_vertices
    out = {"workload": "c5: closure from %d entry points on %d V / %d E kept (%d drawn), one GPU; the stated 20 M / "
                       "500 M graph is the 8-GPU partitioned run" % (sg.entry.size, V, g.info["n_edges"], sg.n_edges),
           "generate_s": gen_s, "upload_s": upload_s, "device_bytes": int(g.info["device_bytes"])}
    o = ora.CsrOracle(V, sg.src, sg.dst, sg.entry)
    t0 = time.perf_counter()
    wlvl, wreached, wte = o.closure(0, sg.entry)
    cpu_closure_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    worder = o.analysis_order()
    cpu_order_s = time.perf_counter() - t0
    kms, kt, wall, st, r = timed_calls(eng, lambda: g.closure(dmgraph.OUT, sg.entry), steps)
    assert r[1] == wreached and np.array_equal(r[0], wlvl), "c5: closure levels differ from the oracle"
    te, rows = st["edges_traversed"], st["rows_expanded"]
    levels = int(wlvl[wlvl != 0xFFFFFFFF].max()) + 1
    alg = 8 * rows + 4 * te + (V // 8) * levels + 4 * V                       # SURVEY.md 8d closure bytes
    out["closure"] = {"kernel_ms": kms, "kernels": kt, "reached": int(r[1]), "levels": levels,
                      "edges_traversed": int(te), "edges_read": int(st["edges_read"]),
                      "gteps": te / (kms * 1e-3) / 1e9, "gteps_read": st["edges_read"] / (kms * 1e-3) / 1e9,
                      "roofline": roofline(alg, kms), "e2e": {"ms": wall, "d2h_bytes": int(st["d2h_bytes"])}}
    kms, kt, wall, st, r = timed_calls(eng, lambda: g.analysis_order(), steps)
    assert np.array_equal(r, worder), "c5: analysisOrder differs from the oracle"
    alg = 2 * (8 * rows + 4 * te) + 4 * V + 4 * V
    out["analysis_order"] = {"kernel_ms": kms, "kernels": kt, "roofline": roofline(alg, kms),
                             "e2e": {"ms": wall, "d2h_bytes": int(st["d2h_bytes"])}}
    out["parity_checked"] = 2 * V
    out["parity_against"] = "oracle csr: every level word and the whole exact order"
    out["cpu_baseline"] = {"kind": "port", "cores": 1, "closure_s": cpu_closure_s, "analysis_order_s": cpu_order_s,
                           "gteps_closure": wte / cpu_closure_s / 1e9,
                           "sample": "the whole graph, integer-CSR restatement of PG:390-421 (single thread, as the reference)"}
    g.release()
    return out


# ---------------------------------------------------------------------------------------------------
def run_dist_closure(n_gpus, vertices=4_000_000, edges=100_000_000, steps=5):
    """The one data-path exchange of the hot path: the vertex-partitioned closure over
    `n_gpus` devices of this box through dmg_dist_* (ONE process drives all devices, as the single-JVM host
    would).  Both transports are timed on the device and checked against the oracle: the persistent
    peer-memory kernel (default) and the per-level NCCL allgather.  bench.py --gpus N (N > 1) calls this on
    rank 0 after the replicas' timed regions; run alone: python bench_configs.py dist --gpus N."""
    from oracle import oracle as ora
    t0 = time.perf_counter()
    sg = synth.big_graph_fast(vertices, edges, cap=2000, entry_frac=0.01, seed=synth.SEED_C5)
    gen_s = time.perf_counter() - t0
    o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry)
    t0 = time.perf_counter()
    wlvl, wreached, wte = o.closure(0, sg.entry)
    cpu_s = time.perf_counter() - t0
    dist = dmgraph.Dist(n_gpus)
    t0 = time.perf_counter()
    pg = dmgraph.DistGraph(dist, sg.n_vertices, sg.src, sg.dst, sg.entry, partitioned=True)
    upload_s = time.perf_counter() - t0
    V = sg.n_vertices
    out = {"workload": "partitioned closure from %d entry points, %d V / %d edges drawn, vertex ranges over %d GPUs "
                       "(dmg_dist_closure, one process)" % (sg.entry.size, V, sg.n_edges, n_gpus),
           "n_gpus": n_gpus, "reached": int(wreached), "definitional_edges": int(wte), "generate_s": gen_s,
           "upload_s": upload_s, "parity_against": "oracle csr: every level word, both transports",
           "parity_checked": 2 * V,
           "cpu_baseline": {"kind": "port", "cores": 1, "closure_s": cpu_s, "gteps": wte / cpu_s / 1e9}}
    try:
        for name, env in (("peer_memory", None), ("nccl_allgather", "1")):
            if env:
                os.environ["DMG_DIST_NCCL"] = env
            else:
                os.environ.pop("DMG_DIST_NCCL", None)
            lvl, reached, st = pg.closure(dmgraph.OUT, sg.entry)           # warm-up + check
            assert reached == wreached and np.array_equal(lvl, wlvl), "dist closure (%s) differs from the oracle" % name
            ms = []
            for _ in range(steps):
                lvl, reached, st = pg.closure(dmgraph.OUT, sg.entry)
                ms.append(st["device_ms"])
            assert reached == wreached and np.array_equal(lvl, wlvl), "dist closure (%s) differs from the oracle" % name
            best = float(min(ms))
            levels = int(st["levels"])
            alg = 8 * int(wreached) + 4 * int(wte) + (V // 8) * levels + 4 * V   # SURVEY 8d closure bytes, whole job
            out[name] = {"device_ms": best, "device_ms_median": float(np.median(ms)), "levels": levels,
                         "push_levels": int(st["push_levels"]), "pull_levels": int(st["pull_levels"]),
                         "transport": int(st["transport"]), "edges_read": int(st["edges_read"]),
                         "exchanged_bytes_per_rank": int(st["exchanged_bytes_per_rank"]),
                         "gteps_definitional": wte / (best * 1e-3) / 1e9,
                         "roofline": roofline(alg, best, peak_all_gpus=HBM * n_gpus,
                                              frac_all_gpus=alg / (best * 1e-3) / 1e9 / (HBM * n_gpus))}
    finally:
        os.environ.pop("DMG_DIST_NCCL", None)
        pg.release()
        dist.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["c3", "c4", "c5", "dist"])
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--names", type=int, default=10_000_000)
    ap.add_argument("--queries", type=int, default=16384)
    ap.add_argument("--vertices", type=int, default=None)
    ap.add_argument("--edges", type=int, default=None)
    ap.add_argument("--ops", type=int, default=262_144)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    if args.config == "dist":
        line = run_dist_closure(args.gpus, args.vertices or 4_000_000, args.edges or 100_000_000, args.steps)
        line["config"] = "dist"
        print(json.dumps(line), flush=True)
        return
    eng = dmgraph.Engine(0)
    if args.config == "c3":
        line = run_c3(eng, args.names, args.queries, args.steps)
    elif args.config == "c4":
        line = run_c4(eng, args.vertices or 1_000_000, args.edges or 30_000_000, n_ops=args.ops, steps=args.steps)
    else:
        line = run_c5(eng, args.vertices or 4_000_000, args.edges or 100_000_000, args.steps)
    line["config"] = args.config
    print(json.dumps(line), flush=True)
    eng.close()


if __name__ == "__main__":
    main()
