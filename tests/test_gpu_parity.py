"""GPU parity: the CUDA path, called through the C ABI, against the oracle and the golden
vectors.  Bit-exact: ids, hops and orders are integers; nothing here has a tolerance."""
import numpy as np
import pytest

import dmgraph
import golden_runner as gr
from dmgraph import synth
from engine_adapter import EngineGraph
from oracle import oracle as ora
from oracle import pyref

pytestmark = pytest.mark.gpu

CASES = gr.load_cases()


@pytest.fixture(scope="module")
def engine():
    e = dmgraph.Engine(0)
    yield e
    e.close()


# ---- golden vectors transcribed from the reference's unit tests ------------------------------
@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_reference_unit_tests(engine, case):
    g = gr.build(case, EngineGraph(engine))
    gr.check(case, g)


# ---- BASELINE config 1: the reference indexing itself, 2-hop for every class ----------------
def _self_index(engine):
    d = gr.load_self_index()
    g, p = EngineGraph(engine), pyref.ProjectGraph()
    for i, f in d["nodes"]:
        g.add_node(i, f)
        p.add_node(i, f)
    for a, b in d["edges"]:
        g.add_dependency(a, b)
        p.add_dependency(a, b)
    for e in d["entryPoints"]:
        g.mark_as_entry_point(e)
        p.mark_as_entry_point(e)
    return g, p


def test_config1_self_index(engine):
    g, p = _self_index(engine)
    names = p.identifiers()
    assert g.node_count() == 46 and g.graph().info["n_edges"] == 77
    assert g.analysis_order() == p.analysis_order()
    assert g.entry_points() == p.entry_points()
    n = len(names)
    q_off = np.arange(n + 1, dtype=np.uint32)
    q_src = np.arange(n, dtype=np.uint32)
    for mode in ("out", "in", "both"):
        for k in (1, 2, 3):
            for excl in (True, False):
                off, ids, hop = g.graph().neighbors_batch(dmgraph.MODES[mode], k, excl, q_off, q_src)
                for v, nm in enumerate(names):
                    want = pyref.khop(p, [nm], k, mode, excl)
                    a, b = int(off[v]), int(off[v + 1])
                    assert {names[i]: int(h) for i, h in zip(ids[a:b], hop[a:b])} == want
                    assert list(ids[a:b]) == sorted(ids[a:b])
    for nm in names:
        assert g.dependencies(nm) == p.dependencies(nm)                      # exact order
        assert g.dependents(nm) == sorted(p.dependents(nm), key=names.index)  # canonical order
        assert sorted(g.resolve(nm)) == sorted(p.resolve(nm))
    for q in ["Graph", "service", "", "co.fanki", "zzz", "Controller", "r", "domainmcp.shared.P"]:
        assert g.search_project(q) == pyref.search_project(p, q)
    for t in ["GraphService", "graphservice", "Preconditions", "Nope", "Parser", "shared"]:
        assert g.resolve_class_name(t) == pyref.resolve_class_name(p, t)


# ---- differential tests on seeded random graphs ----------------------------------------------
def _random_graph(rng, n, e, n_parts=1, dup=True):
    """Edges inside random project ranges, with duplicates, self-loops and dangling endpoints."""
    cuts = np.sort(rng.choice(np.arange(1, n), size=min(n_parts - 1, max(n - 1, 0)), replace=False)) if n_parts > 1 else []
    part_off = np.concatenate([[0], cuts, [n]]).astype(np.uint32)
    part_of = np.searchsorted(part_off, np.arange(n), side="right") - 1
    src = rng.integers(0, n, e)
    lo, hi = part_off[part_of[src]], part_off[part_of[src] + 1]
    dst = (lo + (rng.random(e) * (hi - lo)).astype(np.int64)).astype(np.int64)
    if dup and e > 4:
        k = e // 8
        src[:k], dst[:k] = src[-k:], dst[-k:]                       # duplicates
        bad = rng.integers(0, e, max(1, e // 20))
        dst[bad] = n + rng.integers(0, 5, bad.size)                 # dangling
    entry = rng.integers(0, n + 2, int(rng.integers(0, 6)))
    return src.astype(np.uint32), dst.astype(np.uint32), entry.astype(np.uint32), part_off


def _random_queries(rng, n, nq, max_src):
    cnt = rng.integers(0, max_src + 1, nq)
    q_off = np.zeros(nq + 1, np.uint32)
    np.cumsum(cnt, out=q_off[1:])
    q_src = rng.integers(0, n + 3, int(q_off[-1])).astype(np.uint32)   # some ids out of range
    return q_off, q_src


@pytest.mark.parametrize("n_parts", [1, 7])
def test_khop_random_small_windows(engine, n_parts):
    rng = np.random.default_rng(100 + n_parts)
    for trial in range(12):
        n = int(rng.integers(2, 600))
        e = int(rng.integers(0, 8 * n))
        src, dst, entry, part_off = _random_graph(rng, n, e, n_parts)
        g = dmgraph.Graph(engine, n, src, dst, entry, part_off=part_off)
        o = ora.CsrOracle(n, src, dst, entry, part_off)
        assert g.info["n_edges"] == o.n_edges
        q_off, q_src = _random_queries(rng, n, 200, 6)
        for mode in (0, 1, 2):
            for k in (1, 2, 5):
                for excl in (0, 1):
                    off, ids, hop = g.neighbors_batch(mode, k, excl, q_off, q_src)
                    woff, wids, whop, wte = o.khop_batch(mode, k, excl, q_off, q_src)
                    assert np.array_equal(off, woff)
                    assert np.array_equal(ids, wids)
                    assert np.array_equal(hop, whop)
                    assert engine.last_stats()["edges_traversed"] == wte
        g.release()


@pytest.mark.parametrize("n", [2100, 4000, 4200, 8192])
def test_khop_random_mid_windows(engine, n):
    """Projects of 2049..8192 vertices: the 4096 and 8192 shared-memory classes of khop_small."""
    rng = np.random.default_rng(n)
    src, dst, entry, part_off = _random_graph(rng, n, 6 * n, 1)
    g = dmgraph.Graph(engine, n, src, dst, entry, part_off=part_off)
    o = ora.CsrOracle(n, src, dst, entry, part_off)
    q_off, q_src = _random_queries(rng, n, 150, 8)
    for mode, k, excl in [(0, 1, 0), (1, 2, 1), (2, 1, 1), (2, 2, 0), (2, 3, 1), (0, 5, 1)]:
        off, ids, hop = g.neighbors_batch(mode, k, excl, q_off, q_src)
        woff, wids, whop, wte = o.khop_batch(mode, k, excl, q_off, q_src, threads=8)
        assert np.array_equal(off, woff)
        assert np.array_equal(ids, wids)
        assert np.array_equal(hop, whop)
        assert engine.last_stats()["edges_traversed"] == wte
    _stride, win, rows = g.neighbors_batch_rows(2, 2, 1, q_off, q_src)
    woff, wids, whop, _ = o.khop_batch(2, 2, 1, q_off, q_src, threads=8)
    for q in range(q_off.size - 1):
        a, b = int(woff[q]), int(woff[q + 1])
        r = rows[q]
        got = np.flatnonzero((r != 0) & (r != 255))
        assert np.array_equal(got + int(win[q, 0]), wids[a:b]) and np.array_equal(r[got], whop[a:b])
    g.release()


def test_khop_large_batch(engine):
    """A batch of several waves of CTAs over many projects, an unknown frame, and the ECAP answer."""
    sg = synth.eco_graph(n_projects=8, v_per=600, e_per=5000, cap=80, seed=31)
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off)
    o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry, sg.part_off)
    q_off, q_src = synth.stacktrace_queries(sg, n_queries=21000, frames=6, seed=32)
    q_src[17] = 0xFFFFFFF0
    for mode, k, excl in [(2, 2, 1), (0, 2, 0), (2, 1, 1)]:
        off, ids, hop = g.neighbors_batch(mode, k, excl, q_off, q_src)
        woff, wids, whop, wte = o.khop_batch(mode, k, excl, q_off, q_src, threads=8)
        assert np.array_equal(off, woff) and np.array_equal(ids, wids) and np.array_equal(hop, whop)
        assert engine.last_stats()["edges_traversed"] == wte
        # a large batch on small projects crosses PCIe as bit rows (expanded by the library's host threads): far
        # fewer bytes than the 5 per entry of the packed rows, but never none
        assert 0 < engine.last_stats()["d2h_bytes"] < 5 * ids.size
    with pytest.raises(dmgraph.DmgError) as ei:
        g.neighbors_batch(2, 2, 1, q_off, q_src, cap=int(woff[-1]) // 2)
    assert ei.value.code == dmgraph.ECAP
    g.release()


def test_khop_random_large_window(engine):
    """Projects above 8192 vertices take the global-scratch kernel, incl. its long-row path."""
    rng = np.random.default_rng(5)
    for n, e, n_parts in [(9000, 30000, 1), (20000, 150000, 1), (30000, 60000, 2)]:
        src, dst, entry, part_off = _random_graph(rng, n, e, n_parts)
        g = dmgraph.Graph(engine, n, src, dst, entry, part_off=part_off)
        o = ora.CsrOracle(n, src, dst, entry, part_off)
        q_off, q_src = _random_queries(rng, n, 300, 4)
        for mode, k, excl in [(0, 1, 0), (1, 1, 0), (2, 1, 1), (2, 2, 1), (0, 4, 0), (2, 6, 1), (1, 3, 1)]:
            off, ids, hop = g.neighbors_batch(mode, k, excl, q_off, q_src)
            woff, wids, whop, wte = o.khop_batch(mode, k, excl, q_off, q_src, threads=8)
            assert np.array_equal(off, woff)
            assert np.array_equal(ids, wids)
            assert np.array_equal(hop, whop)
            assert engine.last_stats()["edges_traversed"] == wte
        g.release()


def test_adjacency_random(engine):
    rng = np.random.default_rng(9)
    for n, e in [(1, 0), (50, 400), (3000, 40000)]:
        src, dst, entry, part_off = _random_graph(rng, n, e, 1)
        g = dmgraph.Graph(engine, n, src, dst, entry)
        o = ora.CsrOracle(n, src, dst, entry)
        ids = rng.integers(0, n + 5, 1000).astype(np.uint32)
        for d in (0, 1):
            off, out = g.adjacency_batch(d, ids)
            woff, wout = o.adjacency_batch(d, ids)
            assert np.array_equal(off, woff) and np.array_equal(out, wout)
        # all-vertex dump (classes:+dependencies:+dependents, GQS:111-150)
        allv = np.arange(n, dtype=np.uint32)
        for d in (0, 1):
            off, out = g.adjacency_batch(d, allv)
            woff, wout = o.adjacency_batch(d, allv)
            assert np.array_equal(off, woff) and np.array_equal(out, wout)
        assert np.array_equal(g.entry_points(), o.entry_points())
        flags = g.is_entry_point_batch(ids)
        es = set(o.entry_points().tolist())
        assert flags.tolist() == [int(i in es) for i in ids.tolist()]
        g.release()


def test_adjacency_hub_rows(engine):
    """Zipf hubs: rows far above the per-warp copy limit go through the grid-wide copy."""
    n = 9000
    leaves = np.arange(1, n, dtype=np.uint32)
    src = np.concatenate([leaves, np.zeros(3000, np.uint32), leaves[:4000]])
    dst = np.concatenate([np.zeros(n - 1, np.uint32), np.arange(1, 3001, dtype=np.uint32), np.full(4000, 7, np.uint32)])
    g = dmgraph.Graph(engine, n, src, dst, [])
    o = ora.CsrOracle(n, src, dst, [])
    ids = np.array([0, 7, 5, 0, 8999, 7, 9000, 1], np.uint32)
    for d in (0, 1):
        off, out = g.adjacency_batch(d, ids)
        woff, wout = o.adjacency_batch(d, ids)
        assert np.array_equal(off, woff) and np.array_equal(out, wout)
    assert int(g.adjacency_batch(1, [0])[0][1]) == n - 1
    g.release()


def test_empty_and_ragged_inputs(engine):
    g = dmgraph.Graph(engine, 0, [], [], [])
    off, ids, hop = g.neighbors_batch(2, 1, 1, [0, 0, 2], [0, 5])
    assert off.tolist() == [0, 0, 0] and ids.size == 0
    assert g.analysis_order().size == 0
    lvl, reached = g.closure(0, [])
    assert reached == 0
    g.release()
    g = dmgraph.Graph(engine, 3, [0, 1], [1, 2], [0])
    off, ids, hop = g.neighbors_batch(2, 1, 1, [0], [])            # zero queries
    assert off.tolist() == [0]
    off, ids, hop = g.neighbors_batch(2, 2, 1, [0, 0, 1, 1], [1])  # empty rows around a real one
    assert off.tolist() == [0, 0, 2, 2] and ids.tolist() == [0, 2] and hop.tolist() == [1, 1]
    with pytest.raises(dmgraph.DmgError) as ei:
        g.neighbors_batch(2, 0, 1, [0, 1], [0])
    assert ei.value.code == dmgraph.EINVAL
    with pytest.raises(dmgraph.DmgError) as ei:
        g.neighbors_batch(2, 1, 1, [0, 1], [1], cap=1)             # needs 2
    assert ei.value.code == dmgraph.ECAP
    with pytest.raises(dmgraph.DmgError):                           # edge across project ranges
        dmgraph.Graph(engine, 4, [0], [3], [], part_off=[0, 2, 4])
    g.release()


# ---- closure / analysisOrder ------------------------------------------------------------------
def test_closure_and_order_random(engine):
    rng = np.random.default_rng(21)
    for n, e in [(2, 1), (40, 60), (500, 900), (5000, 20000), (20000, 30000), (40000, 90000), (150000, 900000)]:
        src, dst, entry, part_off = _random_graph(rng, n, e, 1)
        g = dmgraph.Graph(engine, n, src, dst, entry)
        o = ora.CsrOracle(n, src, dst, entry)
        assert np.array_equal(g.analysis_order(), o.analysis_order())
        # the BFS part of analysisOrder is the OUT closure of the entry points: same TE
        assert engine.last_stats()["edges_traversed"] == o.closure(0, entry)[2]
        for d in (0, 1):
            s = rng.integers(0, n + 2, int(rng.integers(0, 5))).astype(np.uint32)
            lvl, reached = g.closure(d, s)
            wl, wr, wte = o.closure(d, s)
            assert np.array_equal(lvl, wl) and reached == wr
            assert engine.last_stats()["edges_traversed"] == wte
        g.release()


def test_order_deep_chain(engine):
    n = 3000
    src = np.arange(n - 1)
    g = dmgraph.Graph(engine, n, src + 1, src, [n - 1])     # n-1 -> n-2 -> ... -> 0
    assert g.analysis_order().tolist() == list(range(n - 1, -1, -1))
    lvl, reached = g.closure(0, [n - 1])
    assert reached == n and lvl.tolist() == list(range(n - 1, -1, -1))
    g.release()


def test_closure_hybrid_shapes(engine):
    """>= 32768 vertices take the bitmap push/pull closure: a deep chain (push only), a star and a
    dense random graph (pull levels), all four direction / source combinations."""
    n = 40000
    chain = np.arange(n - 1, dtype=np.uint32)
    g = dmgraph.Graph(engine, n, chain, chain + 1, [])
    lvl, reached = g.closure(0, [0])
    assert reached == n and np.array_equal(lvl, np.arange(n, dtype=np.uint32))
    lvl, reached = g.closure(1, [n - 1, n + 5])
    assert reached == n and np.array_equal(lvl, np.arange(n - 1, -1, -1, dtype=np.uint32))
    g.release()
    rng = np.random.default_rng(8)
    src = rng.integers(0, n, 600000).astype(np.uint32)
    dst = rng.integers(0, n, 600000).astype(np.uint32)
    g = dmgraph.Graph(engine, n, src, dst, [])
    o = ora.CsrOracle(n, src, dst, [])
    for d in (0, 1):
        for s in ([7], list(range(0, n, 3))):
            lvl, reached = g.closure(d, s)
            wl, wr, wte = o.closure(d, s)
            assert reached == wr and np.array_equal(lvl, wl)
            st = engine.last_stats()
            assert st["edges_traversed"] == wte and st["edges_read"] <= 2 * wte + n
    g.release()


# ---- search -----------------------------------------------------------------------------------
def test_search_random(engine):
    rng = np.random.default_rng(33)
    names = synth.project_names(3, 3000, rng)
    off, arena = synth.pack_names(names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    g = dmgraph.Graph(engine, len(names), [], [], [], name_off=off, name_bytes=arena)
    queries = [b"", b"order", b"svc3", b"x", b"controller", b"zzzz", b"1", b".", b"payment12",
               names[7].lower().encode(), names[2999].lower().encode()[-9:], b"o" * 2000]
    queries += [names[int(i)].lower().encode()[int(a):int(a) + int(l)]
                for i, a, l in zip(rng.integers(0, 3000, 40), rng.integers(0, 20, 40), rng.integers(1, 12, 40))]
    for kind in (0, 1, 2):
        for (lo, hi) in [(0, None), (100, 1700), (2999, 3000), (5, 5)]:
            o, ids = g.search_batch(kind, queries, lo, hi)
            wo, wids = ora.search_batch(off, low, queries, kind, lo, hi)
            assert np.array_equal(o, wo) and np.array_equal(ids, wids)
    g.release()


# ---- full-size properties (BASELINE config 2 shape) -------------------------------------------
def test_config2_full_size_properties(engine):
    sg = synth.eco_graph()
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off)
    q_off, q_src = synth.stacktrace_queries(sg, n_queries=65536)
    off, ids, hop = g.neighbors_batch(dmgraph.BOTH, 2, 1, q_off, q_src)
    st = engine.last_stats()
    nq = q_off.size - 1
    # rows sorted, inside the query's project, never a source, hops in {1,2}
    row = np.repeat(np.arange(nq), np.diff(off).astype(np.int64))
    assert np.all((hop == 1) | (hop == 2))
    part_of_row = (q_src[q_off[:-1]] // 2000)[row]
    assert np.array_equal(ids // 2000, part_of_row)
    same_row = row[1:] == row[:-1]
    assert np.all(ids[1:][same_row] > ids[:-1][same_row])
    # idempotence
    off2, ids2, hop2 = g.neighbors_batch(dmgraph.BOTH, 2, 1, q_off, q_src)
    assert np.array_equal(off, off2) and np.array_equal(ids, ids2) and np.array_equal(hop, hop2)
    # k=1 is a subset with the same hop-1 set
    off1, ids1, hop1 = g.neighbors_batch(dmgraph.BOTH, 1, 1, q_off, q_src)
    assert np.all(hop1 == 1) and int(off1[-1]) == int((hop == 1).sum())
    assert np.array_equal(ids1, ids[hop == 1])
    # the oracle on a slice of the batch
    o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry, sg.part_off)
    sl = slice(0, 2048)
    woff, wids, whop, wte = o.khop_batch(2, 2, 1, q_off[:2049] - q_off[0], q_src[:int(q_off[2048])], threads=8)
    assert np.array_equal(off[:2049], woff)
    assert np.array_equal(ids[:int(woff[-1])], wids) and np.array_equal(hop[:int(woff[-1])], whop)
    assert st["results"] == int(off[-1])
    g.release()


# ---- GraphService lifecycle (GS:55-122) --------------------------------------------------------
def test_graph_service_put_reload_swap(engine):
    from dmgraph.host import GraphService, ProjectGraph
    svc = GraphService(engine)
    a = ProjectGraph()
    a.add_node("co.fanki.A", "A.java")
    a.add_node("co.fanki.B", "B.java")
    a.add_dependency("co.fanki.A", "co.fanki.B")
    svc.put("p1", "order-service", a)
    assert svc.size() == 1 and svc.get_graph("nope") is None and svc.get_graph_by_project_name("x") is None
    g1 = svc.get_graph_by_project_name("order-service")
    assert g1.adjacency_batch(dmgraph.OUT, [0])[1].tolist() == [1]
    # reload from graph_data JSON swaps the whole graph; the reader's retained handle stays valid
    doc = '{"nodes":{"co.fanki.B":{"sourceFile":"B.java"},"co.fanki.A":{"sourceFile":"A.java"}},' \
          '"edges":{"co.fanki.B":["co.fanki.A","co.fanki.Gone"]},"entryPoints":["co.fanki.B"]}'
    assert svc.load_json("p1", "order-service", doc)
    g2 = svc.get_graph("p1")
    assert g2.info["dropped_edges"] == 1 and g2.adjacency_batch(dmgraph.OUT, [0])[1].tolist() == [1]
    assert g2.entry_points().tolist() == [0]
    assert g1.adjacency_batch(dmgraph.IN, [1])[1].tolist() == [0]      # old graph still alive
    assert not svc.load_json("p2", "broken", '{"nodes":')               # logged-and-skipped analogue
    assert svc.size() == 1
    g1.release()
    g2.release()
    svc.close()


def test_search_multi_pattern_batches(engine):
    """S1 through the single-pass multi-pattern kernel: every length class, duplicate queries,
    patterns that occur several times in one name, very long names (exhaustive fallback),
    multi-byte UTF-8, ranges; ordered hits bit-exact against the oracle."""
    rng = np.random.default_rng(77)
    names = synth.project_names(1, 20000, rng)
    names[5] = "aaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaa"                 # same pattern at many positions
    names[6] = "abcabcabcabc" * 40                                       # long: several 116-byte chunks
    names[7] = "x" * 700 + "needle" + "y" * 300
    names[11] = "co.fanki.été.Über漢字Service"    # non-ASCII bytes pass through
    off, arena = synth.pack_names(names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    g = dmgraph.Graph(engine, len(names), [], [], [], name_off=off, name_bytes=arena, lower_bytes=low)
    lows = [bytes(low[int(off[i]):int(off[i + 1])]) for i in range(len(names))]
    queries = [b"", b"a", b"aa", b"aaa", b"aaaa", b"aaaaaaa", b"aaaaaaaa", b"aaaaaaaaa", b"abcabca", b"bca",
               b"needle", b"xneedley", b"x" * 700, b"y" * 301, b"zzzzzzzzzzzz", b".", b"co", b"svc1", b"order",
               "été".encode(), "漢".encode(), b"controller", b"er", b"1", b"12", b"123"]
    for i, a, l in zip(rng.integers(0, 20000, 300), rng.integers(0, 30, 300), rng.integers(1, 24, 300)):
        queries.append(lows[int(i)][int(a):int(a) + int(l)])
    queries += queries[:40]                                              # duplicates get their own rows
    for kind in (dmgraph.SEARCH_CONTAINS_ALL, dmgraph.SEARCH_CONTAINS_FIRST):
        for lo, hi in [(0, None), (3, 9), (1000, 17000), (19999, 20000), (8, 8)]:
            o, ids = g.search_batch(kind, queries, lo, hi)
            wo, wids = ora.search_batch(off, low, queries, kind, lo, hi, threads=8)
            assert np.array_equal(o, wo) and np.array_equal(ids, wids)
    # capacity protocol
    with pytest.raises(dmgraph.DmgError) as ei:
        g.search_batch(dmgraph.SEARCH_CONTAINS_ALL, [b"svc1"], cap=5)
    assert ei.value.code == dmgraph.ECAP
    g.release()


def test_search_config3_shape_scaled(engine):
    """BASELINE config 3's generator at 1/50 scale (200 k names, 1024 queries) against the oracle."""
    off, arena = synth.catalog_names(200000)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    queries = synth.search_queries(off, arena, 1024)
    g = dmgraph.Graph(engine, off.size - 1, [], [], [], name_off=off, name_bytes=arena, lower_bytes=low)
    o, ids = g.search_batch(dmgraph.SEARCH_CONTAINS_ALL, queries)
    wo, wids = ora.search_batch(off, low, queries, 0, threads=8)
    assert np.array_equal(o, wo) and np.array_equal(ids, wids)
    assert int(o[-1]) > 1024
    g.release()


# ---- vertex-range partitioned closure, two ranks emulated on one GPU ----------------
@pytest.mark.parametrize("direction", [0, 1])
def test_partitioned_closure_two_ranks_lockstep(engine, direction):
    import torch
    from dmgraph import dist as ddist
    rng = np.random.default_rng(40 + direction)
    n, e = 5000, 14000
    src, dst, entry, _ = _random_graph(rng, n, e, 1)
    sources = rng.integers(0, n, 4).astype(np.uint32)
    world = 2
    ranks = []
    for r in range(world):
        wr, v_lo, v_hi = ddist.partition(n, r, world)
        s_own, d_own = ddist.owned_edges(src, dst, v_lo, v_hi)
        g = dmgraph.Graph(engine, n, s_own, d_own, [])
        pc = ddist.PartitionedClosure(n, r, world, ddist.GpuStepper(g, v_lo, v_hi), "cuda", None)
        ranks.append((g, pc))
    words = ranks[0][1].words
    bits = np.zeros(words, np.uint32)
    s = np.unique(sources.astype(np.int64))
    np.bitwise_or.at(bits, s >> 5, (1 << (s & 31)).astype(np.uint32))
    for g, pc in ranks:
        t = torch.from_numpy(bits.view(np.int32)).cuda()
        pc.frontier.copy_(t)
        pc.visited.copy_(t)
        own = s[(s >= pc.v_lo) & (s < pc.v_hi)] - pc.v_lo
        if own.size:
            pc.level[torch.from_numpy(own).cuda()] = 0
    lvl, n_front, reached = 1, int(s.size), int(s.size)
    while n_front > 0:
        news = []
        if lvl % 2 == 1:                                                   # push level
            for g, pc in ranks:
                pc.stepper.push(direction, pc.frontier, pc.contrib)
            gathered = torch.cat([pc.contrib for _g, pc in ranks])
            for g, pc in ranks:
                news.append(pc.stepper.merge(lvl, gathered, world, words, pc.visited, pc.next, pc.level))
        else:                                                              # pull level
            for g, pc in ranks:
                pc.stepper.pull(direction, pc.frontier, pc.visited, pc.slice)
            gathered = torch.cat([pc.slice for _g, pc in ranks])
            for g, pc in ranks:
                news.append(pc.stepper.merge(lvl, gathered, 1, words, pc.visited, pc.next, pc.level))
        assert len(set(news)) == 1
        for g, pc in ranks:
            pc.frontier, pc.next = pc.next, pc.frontier
        n_front = news[0]
        reached += n_front
        lvl += 1
    want, wreached, _ = ora.CsrOracle(n, src, dst).closure(direction, sources)
    level = np.concatenate([pc.level[: pc.v_hi - pc.v_lo].cpu().numpy().view(np.uint32) for _g, pc in ranks])
    assert reached == wreached and np.array_equal(level, want)
    # the single-rank class end to end
    g1 = dmgraph.Graph(engine, n, src, dst, [])
    pc1 = ddist.PartitionedClosure(n, 0, 1, ddist.GpuStepper(g1, 0, n), "cuda", None)
    lv1, r1 = pc1.run(sources, direction)
    assert r1 == wreached and np.array_equal(lv1.cpu().numpy().view(np.uint32), want)
    # the stream-ordered primitives (what the multi-GPU runs use)
    pc2 = ddist.PartitionedClosure(n, 0, 1, ddist.GpuStreamStepper(g1, 0, n), "cuda", None)
    lv2, r2 = pc2.run(sources, direction)
    assert r2 == wreached and np.array_equal(lv2.cpu().numpy().view(np.uint32), want) and pc2.stats["edges"] > 0
    for g, _pc in ranks:
        g.release()
    g1.release()


# ---- Q1: graph_query execution (GraphQueryServiceTest.java) -----------------------------------
def _gq_service(engine):
    from engine_adapter import EngineQueryService
    case = next(c for c in CASES if c["name"] == "gq_check_exists")
    eg = gr.build(case, EngineGraph(engine))
    return EngineQueryService(eg, case["methods"])


def test_graph_query_cases_from_reference_tests(engine):
    svc = _gq_service(engine)
    run = lambda q: svc.execute(pyref.parse_graph_query(q)[1])
    assert run("my-project:classes") == ("classes", ["co.fanki.UserController", "co.fanki.UserService",
                                                     "co.fanki.UserRepository", "co.fanki.UserDto"])   # GQT:163-175
    assert run("my-project:entrypoints") == ("entrypoints", ["co.fanki.UserController"])               # GQT:222-236
    assert run("my-project:UserService")[1]["className"] == "co.fanki.UserService"                     # GQT:258-271
    assert run("my-project:co.fanki.UserService")[1]["className"] == "co.fanki.UserService"            # GQT:274-285
    with pytest.raises(LookupError):                                                                    # GQT:288-298
        run("my-project:Unknown")
    assert run("my-project:UserController:dependencies") == ("dependencies", ["co.fanki.UserService"])  # GQT:329-343
    assert run("my-project:UserService:dependents") == ("dependents", ["co.fanki.UserController"])      # GQT:346-358
    assert run("my-project:UserController:method:createUser") == ("method", "createUser")               # GQT:361-376
    with pytest.raises(LookupError):                                                                    # GQT:379-391
        run("my-project:UserController:method:nope")
    ov = run("my-project:UserController")                                                               # GQT:394-411
    assert ov[0] == "class" and ov[1]["dependencies"] == ["co.fanki.UserService"] and ov[1]["dependents"] == []
    assert run("my-project:UserController:?createUser") == ("check", (True, "createUser"))              # GQT:414-428
    assert run("my-project:UserController:?deleteUser") == ("check", (False, None))                     # GQT:431-443
    assert run("my-project:UserController:?getUsers") == ("check", (True, "getUsers"))                  # GQT:446-460
    assert run("my-project:UserController:?GETUSERS") == ("check", (True, "getUsers"))                  # GQS:366-371
    assert run("my-project:usercontroller:DEPENDENCIES")[1] == ["co.fanki.UserService"]                 # case-insensitive
    assert run("my-project:UserController:whatever")[0] == "class"                                      # GQS:238-239


def test_query_batch_mixed_ops_against_oracle(engine):
    """A BASELINE config 4 style mix in ONE call; every row against the restatements."""
    rng = np.random.default_rng(4)
    n, e = 4000, 30000
    src, dst, entry, _ = _random_graph(rng, n, e, 1)
    names = ["p.k%d.Cls%d" % (i % 7, i if i % 50 else i // 50) for i in range(n)]     # repeated simple names
    off, arena = synth.pack_names(names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    g = dmgraph.Graph(engine, n, src, dst, entry, name_off=off, name_bytes=arena)
    meths = [["m%d" % j if (v + j) % 3 else "Handle%d" % (v % 5) for j in range(v % 4)] for v in range(n)]
    g.set_methods(meths)
    catalog = [m for ms in meths for m in ms]
    o = ora.CsrOracle(n, src, dst, entry)
    ops, want = [], []
    for i in range(600):
        kind = i % 7
        v = int(rng.integers(0, n + 2))
        if kind == 0:
            ops.append({"op": dmgraph.OP_DEPENDENCIES, "vertex": v})
            want.append((o.adjacency_batch(0, [v])[1].tolist(), None))
        elif kind == 1:
            ops.append({"op": dmgraph.OP_DEPENDENTS, "vertex": v})
            want.append((o.adjacency_batch(1, [v])[1].tolist(), None))
        elif kind == 2:
            ops.append({"op": dmgraph.OP_OVERVIEW, "vertex": v})
            a, b = o.adjacency_batch(0, [v])[1].tolist(), o.adjacency_batch(1, [v])[1].tolist()
            want.append((a + b, [0] * len(a) + [1] * len(b)))
        elif kind == 3:
            mode, k, ex = int(rng.integers(0, 3)), int(rng.integers(1, 4)), int(rng.integers(0, 2))
            ops.append({"op": dmgraph.OP_KHOP, "vertex": v, "mode": mode, "k": k, "exclude": ex})
            _o, ids, hop, _ = o.khop_batch(mode, k, ex, [0, 1], [v])
            want.append((ids.tolist(), hop.tolist()))
        elif kind == 4:
            ops.append({"op": dmgraph.OP_ENTRYPOINTS})
            want.append((o.entry_points().tolist(), None))
        elif kind == 5:
            t = ["cls%d" % rng.integers(0, 90), "k3.cls", "nosuch", "CLS7".lower(), "p.k1"][int(rng.integers(0, 5))].encode()
            ops.append({"op": dmgraph.OP_RESOLVE_NAME, "text": t})
            _o1, i1 = ora.search_batch(off, low, [t], 1)
            _o2, i2 = ora.search_batch(off, low, [t], 2)
            want.append(([int(i1[0])], [2]) if i1.size else (([int(i2[0])], [3]) if i2.size else ([], [])))
        else:
            t = ["m1", "handle%d" % (v % 5), "zzz", "M2".lower()][int(rng.integers(0, 4))]
            ops.append({"op": dmgraph.OP_METHOD, "vertex": v, "text": t.encode()})
            base = sum(len(ms) for ms in meths[:v]) if v < n else 0
            hit = next((j for j, m in enumerate(meths[v] if v < n else []) if m.lower() == t), None)
            want.append(([base + hit], None) if hit is not None else ([], None))
    flags = [int(m.startswith("Handle")) for m in catalog]
    g.set_method_flags(flags)
    ops.append({"op": dmgraph.OP_ENDPOINTS})
    want.append(([i for i, f in enumerate(flags) if f], None))
    got = g.query_batch(ops)
    for (gi, ga), (wi, wa) in zip(got, want):
        assert gi == wi
        if wa is not None:
            assert ga == wa
    assert catalog[got[6][0][0]].lower() == ops[6]["text"].decode() if got[6][0] else True
    g.release()


# ---- concurrency: many host threads on one handle (SURVEY 8b "Threading"; GS:34-38) ------------
def test_concurrent_readers_one_graph(engine):
    import threading
    sg = synth.eco_graph(n_projects=8, v_per=600, e_per=6000, cap=100, seed=5)
    names = sg.names
    off, arena = synth.pack_names(names)
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off,
                      name_off=off, name_bytes=arena)
    q_off, q_src = synth.stacktrace_queries(sg, n_queries=2000, seed=9)
    want_k = g.neighbors_batch(dmgraph.BOTH, 2, 1, q_off, q_src)
    ids = np.arange(0, sg.n_vertices, 3, dtype=np.uint32)
    want_a = g.adjacency_batch(dmgraph.IN, ids)
    want_s = g.search_batch(dmgraph.SEARCH_CONTAINS_ALL, [b"order", b"svc3", b"x"])
    want_c = g.closure(dmgraph.OUT, sg.entry[:5])
    errors = []

    def worker(kind):
        try:
            for _ in range(6):
                if kind == 0:
                    got = g.neighbors_batch(dmgraph.BOTH, 2, 1, q_off, q_src)
                    ok = all(np.array_equal(a, b) for a, b in zip(got, want_k))
                elif kind == 1:
                    got = g.adjacency_batch(dmgraph.IN, ids)
                    ok = all(np.array_equal(a, b) for a, b in zip(got, want_a))
                elif kind == 2:
                    got = g.search_batch(dmgraph.SEARCH_CONTAINS_ALL, [b"order", b"svc3", b"x"])
                    ok = all(np.array_equal(a, b) for a, b in zip(got, want_s))
                else:
                    got = g.closure(dmgraph.OUT, sg.entry[:5])
                    ok = np.array_equal(got[0], want_c[0]) and got[1] == want_c[1]
                if not ok:
                    errors.append("mismatch in worker %d" % kind)
        except Exception as ex:      # noqa: BLE001
            errors.append(repr(ex))

    threads = [threading.Thread(target=worker, args=(i % 4,)) for i in range(12)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors[:3]
    g.release()


def test_dense_rows_equal_packed_rows(engine):
    """dmg_neighbors_batch_rows carries the same sets and hops as dmg_neighbors_batch."""
    sg = synth.eco_graph(n_projects=5, v_per=900, e_per=7000, cap=100, seed=12)
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off)
    q_off, q_src = synth.stacktrace_queries(sg, n_queries=700, frames=5, seed=13)
    q_src[3] = 0xFFFFFFF0                                   # an unknown frame
    for mode, k, excl in [(2, 2, 1), (0, 3, 0), (1, 1, 1)]:
        off, ids, hop = g.neighbors_batch(mode, k, excl, q_off, q_src)
        stride, win, rows = g.neighbors_batch_rows(mode, k, excl, q_off, q_src)
        assert stride == 1024 and rows.shape == (700, 1024)
        for q in range(700):
            r = rows[q]
            loc = np.nonzero((r != 0) & (r != 255))[0]
            a, b = int(off[q]), int(off[q + 1])
            assert np.array_equal(loc + win[q, 0], ids[a:b]) and np.array_equal(r[loc], hop[a:b])
            assert win[q, 1] in (0, 900) and not r[900:].any()
    g.release()


# ---- name -> id lookups ------------------------------------------------------
def test_lookup_batch_exact_and_simple(engine):
    rng = np.random.default_rng(3)
    parts = [synth.project_names(p, 1500, rng) for p in range(3)]
    parts[1][10] = parts[0][7]                        # the same FQCN in two projects
    parts[2][5] = "co.fanki.X.Dup"
    parts[2][900] = "other.pkg.dup"                   # same lower-cased simple name, later id
    names = [n for p in parts for n in p]
    off, arena = synth.pack_names(names)
    g = dmgraph.Graph(engine, len(names), [], [], [], part_off=[0, 1500, 3000, 4500], name_off=off, name_bytes=arena)
    ids = {}
    for i, n in enumerate(names):
        ids.setdefault(n, i)
    q = [names[i] for i in rng.integers(0, 4500, 300)] + ["nope.Missing", "", names[7].upper(), names[7]]
    got = g.lookup_batch(dmgraph.LOOKUP_EXACT, [x.encode() for x in q])
    want = [ids.get(x, 0xFFFFFFFF) if x in ids and names[ids[x]] == x else 0xFFFFFFFF for x in q]
    assert got.tolist() == want
    # project-scoped: the duplicate FQCN resolves inside the asked project
    assert g.lookup_batch(dmgraph.LOOKUP_EXACT, [names[7].encode()], 1500, 3000).tolist() == [1510]
    assert g.lookup_batch(dmgraph.LOOKUP_EXACT, [names[7].encode()], 3000, 4500).tolist() == [0xFFFFFFFF]
    # simple names: first id in nodes order, case-insensitive through the host's lowering
    low = [n.lower() for n in names]
    simple = lambda s: s[s.rfind(".") + 1:]
    qs = [simple(low[i]) for i in rng.integers(0, 4500, 200)] + ["dup", "zzz", ""]
    got = g.lookup_batch(dmgraph.LOOKUP_SIMPLE_FIRST, [x.encode() for x in qs])
    want = [next((i for i, n in enumerate(low) if simple(n) == x), 0xFFFFFFFF) for x in qs]
    assert got.tolist() == want
    assert g.lookup_batch(dmgraph.LOOKUP_SIMPLE_FIRST, [b"dup"], 3006, 4500).tolist() == [3900]
    # and through dmg_search_batch(SIMPLE_EQ_FIRST), which now probes the same table
    lowb = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    o, i1 = g.search_batch(dmgraph.SEARCH_SIMPLE_EQ_FIRST, [x.encode() for x in qs], 100, 4400)
    wo, wi = ora.search_batch(off, lowb, [x.encode() for x in qs], 1, 100, 4400)
    assert np.array_equal(o, wo) and np.array_equal(i1, wi)
    # heavily repeated simple names (TS kinds) over ranges that cut into projects: the table keeps
    # one head per (project, simple name) and walks a chain of later duplicates
    hot = sorted({simple(n) for n in low if n.startswith("modules.")})[:6] + ["dup"]
    for _ in range(40):
        a, b = sorted(int(x) for x in rng.integers(0, 4501, 2))
        got = g.lookup_batch(dmgraph.LOOKUP_SIMPLE_FIRST, [x.encode() for x in hot], a, b)
        want = [next((i for i in range(a, b) if simple(low[i]) == x), 0xFFFFFFFF) for x in hot]
        assert got.tolist() == want
    g.release()


# ---- binary CSR snapshot ------------------------------------------------------
def test_snapshot_save_load_round_trip(engine, tmp_path):
    sg = synth.eco_graph(n_projects=3, v_per=400, e_per=3000, cap=60, seed=21)
    off, arena = synth.pack_names(sg.names)
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off,
                      name_off=off, name_bytes=arena)
    path = str(tmp_path / "eco.dmgcsr")
    g.save(path)
    r = dmgraph.Graph.load(engine, path)
    assert {k: v for k, v in r.info.items() if k != "device_bytes"} == {k: v for k, v in g.info.items() if k != "device_bytes"}
    q_off, q_src = synth.stacktrace_queries(sg, n_queries=300, seed=2)
    for a, b in zip(g.neighbors_batch(2, 2, 1, q_off, q_src), r.neighbors_batch(2, 2, 1, q_off, q_src)):
        assert np.array_equal(a, b)
    ids = np.arange(sg.n_vertices, dtype=np.uint32)
    for d in (0, 1):
        for a, b in zip(g.adjacency_batch(d, ids), r.adjacency_batch(d, ids)):
            assert np.array_equal(a, b)
    assert np.array_equal(g.analysis_order(), r.analysis_order())
    assert np.array_equal(g.entry_points(), r.entry_points())
    for a, b in zip(g.search_batch(0, [b"order", b"svc1", b""]), r.search_batch(0, [b"order", b"svc1", b""])):
        assert np.array_equal(a, b)
    assert np.array_equal(g.lookup_batch(0, [sg.names[5].encode()]), r.lookup_batch(0, [sg.names[5].encode()]))
    bad = tmp_path / "bad.dmgcsr"
    bad.write_bytes(b"not a snapshot")
    with pytest.raises(dmgraph.DmgError):
        dmgraph.Graph.load(engine, str(bad))
    g.release()
    r.release()


def test_search_batches_larger_than_one_pass(engine):
    """More queries than one pass over the arena handles (16384 distinct patterns): sliced internally."""
    rng = np.random.default_rng(5)
    names = synth.project_names(2, 4000, rng)
    off, arena = synth.pack_names(names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    g = dmgraph.Graph(engine, len(names), [], [], [], name_off=off, name_bytes=arena)
    lows = [bytes(low[int(off[i]):int(off[i + 1])]) for i in range(len(names))]
    queries = [lows[int(i)][int(a):int(a) + int(l)] + (b"" if j % 3 else str(j).encode())
               for j, (i, a, l) in enumerate(zip(rng.integers(0, 4000, 40000), rng.integers(0, 25, 40000),
                                                 rng.integers(2, 14, 40000)))]
    for kind in (dmgraph.SEARCH_CONTAINS_ALL, dmgraph.SEARCH_CONTAINS_FIRST):
        o, ids = g.search_batch(kind, queries)
        wo, wids = ora.search_batch(off, low, queries, kind, threads=8)
        assert np.array_equal(o, wo) and np.array_equal(ids, wids)
    g.release()
