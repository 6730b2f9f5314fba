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
