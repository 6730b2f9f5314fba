"""CPU: pins both restatements (oracle/pyref.py, oracle/oracle.cc) to the reference's own tests
and to each other."""
import numpy as np
import pytest

import golden_runner as gr
from oracle import oracle as ora
from oracle import pyref

CASES = gr.load_cases()


class PyAdapter(pyref.ProjectGraph):
    def search_project(self, q):
        return pyref.search_project(self, q)

    def resolve_neighbors(self, frames):
        return pyref.resolve_neighbors(self, frames)

    def resolve_class_name(self, n):
        return pyref.resolve_class_name(self, n)

    def method_check(self, ident, check):
        return pyref.method_check(self, ident, check)


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_pyref_matches_reference_tests(case):
    g = gr.build(case, PyAdapter())
    gr.check(case, g, roundtrip=_rt)


def _rt(g):
    r = pyref.ProjectGraph.from_json(g.to_json())
    a = PyAdapter()
    a.nodes, a.edges, a.entry = r.nodes, r.edges, r.entry
    return a


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_cxx_ref_matches_reference_tests(case):
    g = gr.build(case, ora.RefGraph())
    gr.check(case, g)


def _csr_from_pyref(g):
    ids = {n: i for i, n in enumerate(g.identifiers())}
    src, dst = [], []
    for s, ts in g.edges.items():
        for t in ts:
            src.append(ids[s])
            dst.append(ids[t])
    return ids, ora.CsrOracle(len(ids), src, dst, [ids[e] for e in g.entry_points()])


def _self_index_pyref():
    d = gr.load_self_index()
    g = PyAdapter()
    for i, f in d["nodes"]:
        g.add_node(i, f)
    for a, b in d["edges"]:
        g.add_dependency(a, b)
    for e in d["entryPoints"]:
        g.mark_as_entry_point(e)
    return d, g


def test_self_index_shape():
    d, g = _self_index_pyref()
    assert g.node_count() == 46 and sum(len(t) for t in g.edges.values()) == 77
    assert max(len(t) for t in g.edges.values()) == 19
    assert len(g.dependents("co.fanki.domainmcp.shared.Preconditions")) == 10


def test_three_restatements_agree_on_self_index():
    """BASELINE config 1: 2-hop OUT / IN / BOTH for every class, pyref == C++ ref == C++ csr."""
    d, g = _self_index_pyref()
    r = ora.RefGraph()
    for i, f in d["nodes"]:
        r.add_node(i, f)
    for a, b in d["edges"]:
        r.add_dependency(a, b)
    for e in d["entryPoints"]:
        r.mark_as_entry_point(e)
    ids, csr = _csr_from_pyref(g)
    names = g.identifiers()
    assert r.analysis_order() == g.analysis_order()
    assert [names[i] for i in csr.analysis_order()] == g.analysis_order()
    for mi, mode in enumerate(["out", "in", "both"]):
        for k in (1, 2, 3):
            for excl in (True, False):
                q_off = np.arange(len(names) + 1, dtype=np.uint32)
                off, oid, hop, _te = csr.khop_batch(mi, k, excl, q_off, np.arange(len(names), dtype=np.uint32))
                for v, n in enumerate(names):
                    want = pyref.khop(g, [n], k, mode, excl)
                    assert r.khop([n], k, mode, excl) == want
                    got = {names[i]: int(h) for i, h in zip(oid[off[v]:off[v + 1]], hop[off[v]:off[v + 1]])}
                    assert got == want
                    assert list(oid[off[v]:off[v + 1]]) == sorted(oid[off[v]:off[v + 1]])
    for n in names:
        assert r.dependencies(n) == g.dependencies(n)
        assert sorted(r.dependents(n)) == sorted(g.dependents(n))
        off, out = csr.adjacency_batch(0, [ids[n]])
        assert [names[i] for i in out] == g.dependencies(n)
        off, out = csr.adjacency_batch(1, [ids[n]])
        assert [names[i] for i in out] == sorted(g.dependents(n), key=lambda s: ids[s])


def test_restatements_agree_on_random_graphs():
    rng = np.random.default_rng(7)
    for trial in range(40):
        n = int(rng.integers(1, 40))
        e = int(rng.integers(0, 4 * n + 1))
        g, r = PyAdapter(), ora.RefGraph()
        names = ["p.k%d.C%d" % (i % 3, i) for i in range(n)]
        for nm in names:
            g.add_node(nm, "f")
            r.add_node(nm, "f")
        src = rng.integers(0, n + 2, e)      # some endpoints out of range -> dropped
        dst = rng.integers(0, n + 2, e)
        for a, b in zip(src, dst):
            an = names[a] if a < n else "x.Unknown%d" % a
            bn = names[b] if b < n else "x.Unknown%d" % b
            g.add_dependency(an, bn)
            r.add_dependency(an, bn)
        ent = rng.integers(0, n, int(rng.integers(0, 4)))
        for x in ent:
            g.mark_as_entry_point(names[x])
            r.mark_as_entry_point(names[x])
        csr = ora.CsrOracle(n, src, dst, ent)
        assert r.analysis_order() == g.analysis_order()
        assert [names[i] for i in csr.analysis_order()] == g.analysis_order()
        lvl, reached, _ = csr.closure(0, ent)
        want = pyref.khop(g, [names[x] for x in ent], 10 ** 6, "out", True)
        for v in range(n):
            if names[v] in want:
                assert lvl[v] == want[names[v]]
            elif v in set(int(x) for x in ent):
                assert lvl[v] == 0
            else:
                assert lvl[v] == 0xFFFFFFFF
        for _ in range(6):
            ns = int(rng.integers(0, 5))
            s = rng.integers(0, n + 1, ns)
            k = int(rng.integers(1, 4))
            mi = int(rng.integers(0, 3))
            excl = bool(rng.integers(0, 2))
            sn = [names[x] for x in s if x < n]
            want = pyref.khop(g, sn, k, ["out", "in", "both"][mi], excl)
            assert r.khop(sn, k, ["out", "in", "both"][mi], excl) == want
            off, oid, hop, _ = csr.khop_batch(mi, k, excl, [0, ns], s)
            assert {names[i]: int(h) for i, h in zip(oid, hop)} == want


def test_search_restatements_agree():
    d, g = _self_index_pyref()
    names = g.identifiers()
    from dmgraph import synth
    off, arena = synth.pack_names(names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    queries = ["Graph", "service", "", "co.fanki", "zzz", "Controller", "r", "domainmcp.shared.P"]
    qb = [q.lower().encode() for q in queries]
    o, ids = ora.search_batch(off, low, qb, kind=0)
    r = ora.RefGraph()
    for i, f in d["nodes"]:
        r.add_node(i, f)
    for qi, q in enumerate(queries):
        want, total = pyref.search_project(g, q)
        assert [names[i] for i in ids[o[qi]:o[qi + 1]]] == want
        assert r.search_project(q)[0] == want
        assert total == 46
    # S2 (resolveClassName) steps 2 and 3 as scans
    for t in ["GraphService", "graphservice", "Preconditions", "Nope", "Parser", "shared"]:
        want = pyref.resolve_class_name(g, t)
        assert r.resolve_class_name(t) == want
        o1, i1 = ora.search_batch(off, low, [t.lower().encode()], kind=1)
        o2, i2 = ora.search_batch(off, low, [t.lower().encode()], kind=2)
        got = names[i1[0]] if i1.size else (names[i2[0]] if i2.size else None)
        assert got == want


# ---- S1 / S2 fuzz: the C++ scans against the one-line Python definitions -------------------------
try:
    from hypothesis import given, settings, strategies as st

    _seg = st.text(alphabet="abAB.x/_1", min_size=0, max_size=6)
    _names = st.lists(st.builds(lambda a, b: (a + "." + b) if a else b, _seg, _seg.filter(lambda s: s != "")),
                      min_size=1, max_size=12)

    @settings(max_examples=150, deadline=None)
    @given(_names, st.lists(st.text(alphabet="abAB.x/_1", max_size=4), min_size=1, max_size=6), st.data())
    def test_search_fuzz_against_definitions(names, extra, data):
        from dmgraph import synth
        off, arena = synth.pack_names(names)
        low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
        lows = [n.lower() for n in names]
        subs = []
        for _ in range(4):                                   # substrings of real names hit; random text mostly misses
            n = data.draw(st.sampled_from(lows))
            a = data.draw(st.integers(0, len(n)))
            b = data.draw(st.integers(a, len(n)))
            subs.append(n[a:b])
        queries = [q.lower() for q in extra] + subs
        qb = [q.encode() for q in queries]
        v_lo = data.draw(st.integers(0, len(names)))
        v_hi = data.draw(st.integers(v_lo, len(names)))
        simple = lambda s: s[s.rfind(".") + 1:]
        for lo, hi in ((0, len(names)), (v_lo, v_hi)):
            o, ids = ora.search_batch(off, low, qb, 0, lo, hi)
            for qi, q in enumerate(queries):                 # CCS:1511-1526: nodes order, "" matches everything
                assert ids[o[qi]:o[qi + 1]].tolist() == [i for i in range(lo, hi) if q in lows[i]]
            o, ids = ora.search_batch(off, low, qb, 1, lo, hi)
            for qi, q in enumerate(queries):                 # GQS:434-446: first simple-name equality
                want = next((i for i in range(lo, hi) if simple(lows[i]) == q), None)
                assert ids[o[qi]:o[qi + 1]].tolist() == ([] if want is None else [want])
            o, ids = ora.search_batch(off, low, qb, 2, lo, hi)
            for qi, q in enumerate(queries):                 # GQS:449-453: first containment
                want = next((i for i in range(lo, hi) if q in lows[i]), None)
                assert ids[o[qi]:o[qi + 1]].tolist() == ([] if want is None else [want])
except ImportError:                                          # pragma: no cover
    pass


def test_multi_project_windows_against_single_project_restatement():
    """One handle over several projects (quirk 6, CCS:1426-1434, 1558): a query runs in
    the project of its first known source, sources of other projects match nothing.  The csr
    restatement with part_off must equal the Python restatement run on that one project."""
    rng = np.random.default_rng(77)
    for trial in range(25):
        n_parts = int(rng.integers(1, 5))
        sizes = rng.integers(1, 12, n_parts)
        part_off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint32)
        V = int(part_off[-1])
        graphs, src, dst = [], [], []
        for p in range(n_parts):
            g = PyAdapter()
            base, n = int(part_off[p]), int(sizes[p])
            for i in range(n):
                g.add_node("v%d" % (base + i), "f")
            for _ in range(int(rng.integers(0, 3 * n + 1))):
                a, b = int(rng.integers(0, n)), int(rng.integers(0, n))
                g.add_dependency("v%d" % (base + a), "v%d" % (base + b))
                src.append(base + a)
                dst.append(base + b)
            graphs.append(g)
        csr = ora.CsrOracle(V, np.array(src, np.uint32), np.array(dst, np.uint32), [], part_off)
        for _ in range(12):
            s = rng.integers(0, V + 2, int(rng.integers(0, 6)))          # ids >= V are unknown frames
            k, mi, excl = int(rng.integers(1, 4)), int(rng.integers(0, 3)), bool(rng.integers(0, 2))
            off, oid, hop, _te = csr.khop_batch(mi, k, excl, [0, s.size], s)
            known = [int(x) for x in s if x < V]
            if not known:
                assert oid.size == 0
                continue
            p = int(np.searchsorted(part_off, known[0], side="right") - 1)
            lo, hi = int(part_off[p]), int(part_off[p + 1])
            inside = ["v%d" % x for x in known if lo <= x < hi]
            want = pyref.khop(graphs[p], inside, k, ["out", "in", "both"][mi], excl)
            assert {"v%d" % i: int(h) for i, h in zip(oid, hop)} == want
            assert oid.tolist() == sorted(oid.tolist())
