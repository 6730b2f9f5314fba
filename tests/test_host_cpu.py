"""CPU: the C++ host-side mirror of ProjectGraph (build rules, wire format) -- no device needed.
Cases follow ProjectGraphTest.java (addNode :26-93, addDependency :98-165, markAsEntryPoint
:170-213, JSON :635-724, 810-826, 958-992, 1145-1192)."""
import json

import pytest

import golden_runner as gr
from dmgraph.host import ProjectGraph
from oracle import pyref


def test_build_rules():
    g = ProjectGraph()
    g.add_node("co.fanki.A", "src/old/A.java")
    g.add_node("co.fanki.B", "src/B.java")
    g.add_node("co.fanki.A", "src/new/A.java")            # keeps position, overwrites file
    assert g.node_count() == 2 and g.identifiers() == ["co.fanki.A", "co.fanki.B"]
    assert g.source_file("co.fanki.A") == "src/new/A.java"
    g.add_dependency("co.fanki.A", "co.fanki.B")
    g.add_dependency("co.fanki.A", "co.fanki.B")          # duplicate collapses
    g.add_dependency("co.fanki.A", "co.fanki.Unknown")    # silently dropped
    g.add_dependency("co.fanki.Unknown", "co.fanki.B")
    g.add_dependency("co.fanki.A", "co.fanki.A")          # self-loop is legal
    assert g.edge_count() == 2
    g.mark_as_entry_point("co.fanki.B")
    g.mark_as_entry_point("co.fanki.B")
    g.mark_as_entry_point("co.fanki.Unknown")
    assert g.entry_point_count() == 1
    assert g.contains("co.fanki.A") and not g.contains("co.fanki.Z") and not g.contains(None)
    assert g.id_of("co.fanki.B") == 1 and g.id_of("nope") == -1


@pytest.mark.parametrize("args", [(None, "f"), ("  ", "f"), ("a", None), ("a", ""), ("\t\n", "f")])
def test_add_node_illegal_argument(args):
    with pytest.raises(ValueError):
        ProjectGraph().add_node(*args)


def test_dependency_and_entry_illegal_argument():
    g = ProjectGraph()
    for a, b in [(None, "co.fanki.B"), ("co.fanki.A", "  "), ("", "x")]:
        with pytest.raises(ValueError):
            g.add_dependency(a, b)
    for a in (None, "", " "):
        with pytest.raises(ValueError):
            g.mark_as_entry_point(a)


def _fill(g, case):
    for op in case["build"]:
        if op[0] == "node":
            g.add_node(op[1], op[2])
        elif op[0] == "dep":
            g.add_dependency(op[1], op[2])
        elif op[0] == "entry":
            g.mark_as_entry_point(op[1])
    return g


@pytest.mark.parametrize("case", [c for c in gr.load_cases() if not any(o[0].endswith("_err") for o in c["build"])],
                         ids=lambda c: c["name"])
def test_json_round_trip_matches_python_restatement(case):
    g = _fill(ProjectGraph(), case)
    p = _fill(pyref.ProjectGraph(), case)
    doc = json.loads(g.to_json())
    assert list(doc["nodes"]) == p.identifiers()
    assert {k: v["sourceFile"] for k, v in doc["nodes"].items()} == p.nodes
    assert {k: list(v) for k, v in doc["edges"].items()} == {k: list(v) for k, v in p.edges.items()}
    assert doc["entryPoints"] == p.entry_points()
    assert doc["methodParameters"] == {} and doc["nodeInfo"] == {} and doc["methodInfo"] == {}
    r = ProjectGraph.from_json(g.to_json())
    assert r.identifiers() == g.identifiers() and r.edge_count() == g.edge_count()
    assert json.loads(r.to_json()) == doc
    # and the reference-format document the Python restatement writes loads identically
    r2 = ProjectGraph.from_json(p.to_json())
    assert json.loads(r2.to_json())["edges"] == doc["edges"]


def test_from_json_old_format_and_metadata_passthrough():
    old = '{"nodes":{"co.fanki.A":{"sourceFile":"src/A.java"}},"edges":{},"entryPoints":[]}'
    g = ProjectGraph.from_json(old)                        # ProjectGraphTest.java:982-992
    assert g.node_count() == 1 and g.edge_count() == 0
    new = ('{"nodes":{"co.fanki.A":{"sourceFile":"src/A.java","classId":"cid-1"},'
           '"co.fanki.\\u00e9B":{"sourceFile":"src/B \\"q\\".java"}},'
           '"edges":{"co.fanki.A":["co.fanki.\\u00e9B","co.fanki.Dangling"]},"entryPoints":["co.fanki.A"],'
           '"methodParameters":{"co.fanki.A":{"m":[{"position":0,"target":"co.fanki.\\u00e9B"}]}},'
           '"nodeInfo":{"co.fanki.A":{"classType":"SERVICE","description":"d {x} [y]"}},'
           '"methodInfo":{"co.fanki.A":[{"methodName":"m","lineNumber":3}]}}')
    g = ProjectGraph.from_json(new)
    assert g.identifiers() == ["co.fanki.A", "co.fanki.éB"]
    assert g.source_file("co.fanki.éB") == 'src/B "q".java'
    assert g.edge_count() == 2                             # dangling kept on the host (PG:782-795)
    assert json.loads(g.to_json()) == json.loads(new)
    for bad in ("", "   ", "[1,2]", '{"nodes":{"a":{}}}', '{"nodes":'):
        with pytest.raises(ValueError):
            ProjectGraph.from_json(bad)


# ---- W0 fuzz: arbitrary reference-shaped documents survive the C++ reader / writer --------------
try:
    from hypothesis import given, settings, strategies as st
    _HAVE_HYPOTHESIS = True
except Exception:                                          # pragma: no cover
    _HAVE_HYPOTHESIS = False

if _HAVE_HYPOTHESIS:
    _name = st.text(alphabet=st.characters(blacklist_categories=("Cs",), min_codepoint=1), min_size=1, max_size=12) \
        .filter(lambda s: s.strip() != "")
    _scalar = st.one_of(st.none(), st.booleans(), st.integers(-2**31, 2**31), st.text(max_size=8))
    _meta = st.recursive(_scalar, lambda c: st.one_of(st.lists(c, max_size=3),
                                                      st.dictionaries(st.text(max_size=6), c, max_size=3)), max_leaves=8)

    @st.composite
    def _documents(draw):
        ids = draw(st.lists(_name, min_size=1, max_size=8, unique=True))
        nodes = {}
        for i in ids:
            n = {"sourceFile": draw(st.text(min_size=1, max_size=10).filter(lambda s: s.strip() != ""))}
            if draw(st.booleans()):
                n["classId"] = draw(st.text(max_size=6))
            nodes[i] = n
        edges = {}
        for s in draw(st.lists(st.sampled_from(ids), max_size=4, unique=True)):
            edges[s] = draw(st.lists(st.sampled_from(ids), max_size=5, unique=True))
        doc = {"nodes": nodes, "edges": edges,
               "entryPoints": draw(st.lists(st.sampled_from(ids), max_size=3, unique=True))}
        if draw(st.booleans()):                            # new-format sections are opaque host payload
            doc["methodParameters"] = draw(st.dictionaries(st.sampled_from(ids), _meta, max_size=2))
            doc["nodeInfo"] = draw(st.dictionaries(st.sampled_from(ids), _meta, max_size=2))
            doc["methodInfo"] = draw(st.dictionaries(st.sampled_from(ids), _meta, max_size=2))
        return doc

    @settings(max_examples=120, deadline=None)
    @given(_documents(), st.booleans())
    def test_json_fuzz_round_trip(doc, ascii_only):
        text = json.dumps(doc, ensure_ascii=ascii_only)    # \\uXXXX escapes (incl. surrogate pairs) or raw UTF-8
        g = ProjectGraph.from_json(text)
        assert g.identifiers() == list(doc["nodes"])
        assert g.edge_count() == sum(len(v) for v in doc["edges"].values())
        back = json.loads(g.to_json())
        assert back["nodes"] == doc["nodes"] and list(back["nodes"]) == list(doc["nodes"])
        assert {k: v for k, v in back["edges"].items() if v} == {k: v for k, v in doc["edges"].items() if v}
        assert back["entryPoints"] == doc["entryPoints"]
        for sec in ("methodParameters", "nodeInfo", "methodInfo"):
            assert back[sec] == doc.get(sec, {})

    @settings(max_examples=150, deadline=None)
    @given(_documents(), st.integers(0, 400), st.sampled_from(["cut", "flip", "drop"]))
    def test_json_fuzz_malformed_never_crashes(doc, pos, how):
        text = json.dumps(doc)
        pos %= max(len(text), 1)
        bad = {"cut": text[:pos], "flip": text[:pos] + "\\x" + text[pos + 1:], "drop": text[:pos] + text[pos + 1:]}[how]
        try:
            json.loads(bad)
            still_json = True
        except Exception:
            still_json = False
        try:
            ProjectGraph.from_json(bad)                    # either loads or raises ValueError; never aborts
        except ValueError:
            assert True
        else:
            assert still_json or how == "drop"             # the reader is no stricter than it needs to be


if _HAVE_HYPOTHESIS:
    _ids = st.sampled_from(["a.A", "a.B", "b.C", "b.c", "x/y", "Z", "é.É", "a.A "])
    _ops = st.lists(st.one_of(
        st.tuples(st.just("node"), _ids, st.sampled_from(["f1", "f2", "dir/f 3"])),
        st.tuples(st.just("dep"), _ids, _ids),
        st.tuples(st.just("entry"), _ids)), max_size=40)

    @settings(max_examples=200, deadline=None)
    @given(_ops)
    def test_builder_fuzz_against_python_restatement(ops):
        """Random addNode / addDependency / markAsEntryPoint sequences (re-adds, unknown endpoints,
        self-loops, duplicates): the C++ mirror and the Python restatement of PG:160-197 agree on
        order, files, edges and entry points, and the JSON round trip keeps all of it."""
        g, p = ProjectGraph(), pyref.ProjectGraph()
        for op in ops:
            for t in (g, p):
                if op[0] == "node":
                    t.add_node(op[1], op[2])
                elif op[0] == "dep":
                    t.add_dependency(op[1], op[2])
                else:
                    t.mark_as_entry_point(op[1])
        assert g.identifiers() == p.identifiers()
        assert g.node_count() == len(p.identifiers()) and g.entry_point_count() == p.entry_point_count()
        assert g.edge_count() == sum(len(v) for v in p.edges.values())
        doc = json.loads(g.to_json())
        assert {k: v["sourceFile"] for k, v in doc["nodes"].items()} == p.nodes
        assert {k: list(v) for k, v in doc["edges"].items() if v} == {k: list(v) for k, v in p.edges.items() if v}
        assert doc["entryPoints"] == p.entry_points()
        for i, name in enumerate(p.identifiers()):
            assert g.id_of(name) == i and g.name_of(i) == name and g.contains(name)
        r = ProjectGraph.from_json(g.to_json())
        assert json.loads(r.to_json()) == doc
