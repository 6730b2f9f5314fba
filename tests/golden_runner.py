"""Replays tests/golden/reference_tests.json against any object that mirrors the reference's
ProjectGraph / CodeContextService / GraphQueryService read API (names in, names out).

"asserted" expectations are the literal assertions of the reference's JUnit tests; "derived_*"
keys hold what the reference's code fixes by construction but no reference test spells out.
Both are checked; the distinction is kept in the fixture for the reader.
"""
import json
import os

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def load_cases():
    with open(os.path.join(HERE, "golden", "reference_tests.json")) as f:
        return json.load(f)["cases"]


def load_self_index():
    with open(os.path.join(HERE, "golden", "self_index.json")) as f:
        return json.load(f)


def build(case, g):
    for op in case["build"]:
        kind = op[0]
        if kind == "node":
            g.add_node(op[1], op[2])
        elif kind == "dep":
            g.add_dependency(op[1], op[2])
        elif kind == "entry":
            g.mark_as_entry_point(op[1])
        elif kind == "node_err":
            with pytest.raises(ValueError):
                g.add_node(op[1], op[2])
        elif kind == "dep_err":
            with pytest.raises(ValueError):
                g.add_dependency(op[1], op[2])
        elif kind == "entry_err":
            with pytest.raises(ValueError):
                g.mark_as_entry_point(op[1])
        else:
            raise AssertionError(kind)
    if "methods" in case and hasattr(g, "methods"):
        g.methods.update(case["methods"])
    return g


def _check_seq(chk, got):
    if "set" in chk:
        assert sorted(got) == sorted(chk["set"]) and len(set(got)) == len(got)
    if "derived_set" in chk:
        assert sorted(got) == sorted(chk["derived_set"])
    if "list" in chk:
        assert list(got) == chk["list"]
    if "derived_list" in chk:
        assert list(got) == chk["derived_list"]
    if "contains" in chk:
        for x in chk["contains"]:
            assert x in got
    if "size" in chk:
        assert len(got) == chk["size"]
    if "count" in chk:
        assert len(got) == chk["count"]
    if "index" in chk:
        for k, v in chk["index"].items():
            assert list(got).index(k) == v
    if "before" in chk:
        for a, b in chk["before"]:
            assert list(got).index(a) < list(got).index(b)
    if "last" in chk:
        assert got[-1] == chk["last"]
    if "first" in chk:
        assert got[0] == chk["first"]
    if chk.get("distinct"):
        assert len(set(got)) == len(got)


def check(case, g, roundtrip=None):
    for chk in case["checks"]:
        op = chk["op"]
        if op == "nodeCount":
            assert g.node_count() == chk["eq"]
        elif op == "contains":
            assert g.contains(chk["arg"]) == chk["eq"]
        elif op == "identifiers":
            _check_seq(chk, g.identifiers())
        elif op == "sourceFile":
            assert g.source_file(chk["arg"]) == chk["eq"]
        elif op == "entryPointCount":
            assert g.entry_point_count() == chk["eq"]
        elif op == "isEntryPoint":
            assert g.is_entry_point(chk["arg"]) == chk["eq"]
        elif op == "entryPoints":
            _check_seq(chk, g.entry_points())
        elif op == "dependencies":
            _check_seq(chk, g.dependencies(chk["arg"]))
        elif op == "dependents":
            _check_seq(chk, g.dependents(chk["arg"]))
        elif op == "resolve":
            _check_seq(chk, g.resolve(chk["arg"]))
        elif op == "analysisOrder":
            _check_seq(chk, g.analysis_order())
        elif op == "jsonRoundTrip":
            if roundtrip is None:
                continue
            r = roundtrip(g)
            for what in chk["same"]:
                if what == "analysisOrder":
                    assert r.analysis_order() == g.analysis_order()
                else:
                    arg = what.split(":", 1)[1]
                    assert sorted(r.resolve(arg)) == sorted(g.resolve(arg))
        elif op == "resolveNeighbors":
            _check_seq(chk, g.resolve_neighbors(chk["frames"]))
        elif op == "classDependencies":
            assert g.dependencies(chk["arg"]) == chk["dependencies"]
            assert sorted(g.dependents(chk["arg"])) == sorted(chk["dependents"])
            assert g.is_entry_point(chk["arg"]) == chk["entryPoint"]
        elif op == "searchProject":
            hits, total = g.search_project(chk["query"])
            assert total == chk["total"]
            _check_seq(chk, hits)
        elif op == "classes":
            _check_seq(chk, g.identifiers())
        elif op == "classesDependencies":
            _check_seq(chk, g.dependencies(chk["arg"]))
        elif op == "entrypoints":
            _check_seq(chk, g.entry_points())
        elif op == "resolveClassName":
            assert g.resolve_class_name(chk["arg"]) == chk["eq"]
        elif op == "vertexDeps":
            ident = g.resolve_class_name(chk["arg"])
            got = g.dependencies(ident) if chk["outgoing"] else g.dependents(ident)
            _check_seq(chk, got)
        elif op == "methodCheck":
            if not hasattr(g, "method_check"):
                continue
            ident = g.resolve_class_name(chk["arg"])
            exists, name = g.method_check(ident, chk["check"])
            assert exists == chk["exists"]
            if exists:
                assert name == chk["methodName"]
        else:
            raise AssertionError("unknown check op " + op)
