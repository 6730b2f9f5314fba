#!/usr/bin/env python3
"""Golden vectors transcribed from the reference's own unit tests.

Every case below restates ONE JUnit test of the reference as data: the graph the
test builds (addNode / addDependency / markAsEntryPoint calls, in order) and what
the test asserts.  Nothing here is computed by our code: the expectations are the
literal values in the reference's assertions ("asserted"), plus -- kept apart
under "derived" -- values the reference's code fixes by construction (e.g. the
LinkedHashMap scan order of search hits, CodeContextService.java:1511) but that
no reference test spells out.

Citations are relative to /root/reference/src/test/java/co/fanki/domainmcp/.

Run:  python tests/golden/transcribe_reference_tests.py   (rewrites reference_tests.json)
"""
import json
import os

PGT = "analysis/domain/ProjectGraphTest.java"
CCT = "analysis/application/CodeContextServiceTest.java"
GQT = "analysis/application/GraphQueryServiceTest.java"
MCT = "config/GraphQueryMcpToolTest.java"
JPT = "analysis/domain/java/JavaSourceParserTest.java"


def node(i, f=None):
    return ["node", i, f if f is not None else "src/" + i.split(".")[-1] + ".java"]


def dep(a, b):
    return ["dep", a, b]


def entry(i):
    return ["entry", i]


# buildSimpleChainGraph(), ProjectGraphTest.java:1221-1231
CHAIN = [node("co.fanki.Controller"), node("co.fanki.Service"), node("co.fanki.Repository"),
         dep("co.fanki.Controller", "co.fanki.Service"), dep("co.fanki.Service", "co.fanki.Repository")]

# buildTestGraph(), GraphQueryServiceTest.java:37-94 (graph part)
GQ_GRAPH = [node("co.fanki.UserController"), node("co.fanki.UserService"),
            node("co.fanki.UserRepository"), node("co.fanki.UserDto"),
            dep("co.fanki.UserController", "co.fanki.UserService"),
            dep("co.fanki.UserService", "co.fanki.UserRepository"),
            entry("co.fanki.UserController")]
GQ_METHODS = {"co.fanki.UserController": ["getUsers", "createUser"],
              "co.fanki.UserService": ["findAll", "create"]}

# buildOrderServiceGraph(), GraphQueryMcpToolTest.java:237-277 (graph part)
MCP_GRAPH = [node("co.fanki.OrderController"), node("co.fanki.OrderService"),
             node("co.fanki.OrderRepository"),
             dep("co.fanki.OrderController", "co.fanki.OrderService"),
             dep("co.fanki.OrderService", "co.fanki.OrderRepository"),
             entry("co.fanki.OrderController")]

CASES = []


def case(name, cite, build, checks, **kw):
    c = {"name": name, "cite": cite, "build": build, "checks": checks}
    c.update(kw)
    CASES.append(c)


# ---- addNode -------------------------------------------------------------------------------
case("addNode_valid", PGT + ":26-33", [node("co.fanki.UserService", "src/UserService.java")],
     [{"op": "nodeCount", "eq": 1}, {"op": "contains", "arg": "co.fanki.UserService", "eq": True}])
case("addNode_multiple", PGT + ":36-49",
     [node("co.fanki.UserService"), node("co.fanki.OrderService"), node("co.fanki.User")],
     [{"op": "nodeCount", "eq": 3},
      {"op": "identifiers", "set": ["co.fanki.UserService", "co.fanki.OrderService", "co.fanki.User"],
       "derived_list": ["co.fanki.UserService", "co.fanki.OrderService", "co.fanki.User"]}])
case("addNode_duplicate_overwrites_file", PGT + ":52-61",
     [node("co.fanki.UserService", "src/old/UserService.java"),
      node("co.fanki.UserService", "src/new/UserService.java")],
     [{"op": "nodeCount", "eq": 1},
      {"op": "sourceFile", "arg": "co.fanki.UserService", "eq": "src/new/UserService.java"}])
case("addNode_null_identifier_throws", PGT + ":64-69", [["node_err", None, "src/A.java"]], [])
case("addNode_blank_identifier_throws", PGT + ":72-77", [["node_err", "  ", "src/A.java"]], [])
case("addNode_null_file_throws", PGT + ":80-85", [["node_err", "co.fanki.A", None]], [])
case("addNode_blank_file_throws", PGT + ":88-93", [["node_err", "co.fanki.A", ""]], [])

# ---- addDependency -------------------------------------------------------------------------
case("addDependency_both_exist", PGT + ":98-107",
     [node("co.fanki.Controller"), node("co.fanki.Service"), dep("co.fanki.Controller", "co.fanki.Service")],
     [{"op": "resolve", "arg": "co.fanki.Controller", "contains": ["co.fanki.Service"]}])
case("addDependency_from_unknown_ignored", PGT + ":110-117",
     [node("co.fanki.Service"), dep("co.fanki.Unknown", "co.fanki.Service")],
     [{"op": "resolve", "arg": "co.fanki.Service", "set": []}])
case("addDependency_to_unknown_ignored", PGT + ":120-127",
     [node("co.fanki.Controller"), dep("co.fanki.Controller", "co.fanki.Unknown")],
     [{"op": "resolve", "arg": "co.fanki.Controller", "set": []}])
case("addDependency_both_unknown_ignored", PGT + ":130-136",
     [dep("co.fanki.A", "co.fanki.B")], [{"op": "nodeCount", "eq": 0}])
case("addDependency_null_from_throws", PGT + ":139-144", [["dep_err", None, "co.fanki.B"]], [])
case("addDependency_blank_to_throws", PGT + ":147-152", [["dep_err", "co.fanki.A", "  "]], [])
case("addDependency_duplicate_collapses", PGT + ":155-165",
     [node("co.fanki.A"), node("co.fanki.B"), dep("co.fanki.A", "co.fanki.B"), dep("co.fanki.A", "co.fanki.B")],
     [{"op": "resolve", "arg": "co.fanki.A", "size": 1}])

# ---- markAsEntryPoint ----------------------------------------------------------------------
case("entry_known", PGT + ":170-177", [node("co.fanki.Controller"), entry("co.fanki.Controller")],
     [{"op": "entryPointCount", "eq": 1}])
case("entry_unknown_ignored", PGT + ":180-186", [entry("co.fanki.Unknown")],
     [{"op": "entryPointCount", "eq": 0}])
case("entry_null_throws", PGT + ":189-194", [["entry_err", None]], [])
case("entry_blank_throws", PGT + ":197-202", [["entry_err", ""]], [])
case("entry_already_marked", PGT + ":205-213",
     [node("co.fanki.Controller"), entry("co.fanki.Controller"), entry("co.fanki.Controller")],
     [{"op": "entryPointCount", "eq": 1}])

# ---- dependencies --------------------------------------------------------------------------
case("dependencies_outgoing", PGT + ":218-225", CHAIN,
     [{"op": "dependencies", "arg": "co.fanki.Controller", "set": ["co.fanki.Service"]}])
case("dependencies_leaf_empty", PGT + ":228-234", CHAIN,
     [{"op": "dependencies", "arg": "co.fanki.Repository", "set": []}])
case("dependencies_unknown_empty", PGT + ":237-241", CHAIN,
     [{"op": "dependencies", "arg": "co.fanki.Unknown", "set": []}])
case("dependencies_null_empty", PGT + ":244-248", CHAIN,
     [{"op": "dependencies", "arg": None, "set": []}])

# ---- dependents ----------------------------------------------------------------------------
case("dependents_incoming", PGT + ":262-269", CHAIN,
     [{"op": "dependents", "arg": "co.fanki.Service", "set": ["co.fanki.Controller"]}])
case("dependents_root_empty", PGT + ":272-278", CHAIN,
     [{"op": "dependents", "arg": "co.fanki.Controller", "set": []}])
case("dependents_leaf", PGT + ":281-287", CHAIN,
     [{"op": "dependents", "arg": "co.fanki.Repository", "set": ["co.fanki.Service"]}])
case("dependents_unknown_empty", PGT + ":290-294", CHAIN,
     [{"op": "dependents", "arg": "co.fanki.Unknown", "set": []}])
case("dependents_null_empty", PGT + ":297-301", CHAIN,
     [{"op": "dependents", "arg": None, "set": []}])

# ---- isEntryPoint / entryPoints ------------------------------------------------------------
case("isEntryPoint_marked", PGT + ":315-321", [node("co.fanki.Controller"), entry("co.fanki.Controller")],
     [{"op": "isEntryPoint", "arg": "co.fanki.Controller", "eq": True}])
case("isEntryPoint_unmarked", PGT + ":324-329", [node("co.fanki.Service")],
     [{"op": "isEntryPoint", "arg": "co.fanki.Service", "eq": False}])
case("isEntryPoint_null", PGT + ":332-336", [], [{"op": "isEntryPoint", "arg": None, "eq": False}])
case("entryPoints_all", PGT + ":339-351",
     [node("co.fanki.A"), node("co.fanki.B"), node("co.fanki.C"), entry("co.fanki.A"), entry("co.fanki.B")],
     [{"op": "entryPoints", "set": ["co.fanki.A", "co.fanki.B"],
       "derived_list": ["co.fanki.A", "co.fanki.B"]}])
case("entryPoints_none", PGT + ":354-359", [node("co.fanki.A")], [{"op": "entryPoints", "set": []}])

# ---- resolve -------------------------------------------------------------------------------
case("resolve_direct", PGT + ":375-381", CHAIN,
     [{"op": "resolve", "arg": "co.fanki.Controller", "contains": ["co.fanki.Service"]}])
case("resolve_reverse", PGT + ":384-391", CHAIN,
     [{"op": "resolve", "arg": "co.fanki.Service", "contains": ["co.fanki.Controller", "co.fanki.Repository"]}])
case("resolve_both", PGT + ":394-401", CHAIN,
     [{"op": "resolve", "arg": "co.fanki.Service", "set": ["co.fanki.Controller", "co.fanki.Repository"]}])
case("resolve_leaf", PGT + ":404-410", CHAIN,
     [{"op": "resolve", "arg": "co.fanki.Repository", "set": ["co.fanki.Service"]}])
case("resolve_null", PGT + ":413-419", CHAIN, [{"op": "resolve", "arg": None, "set": []}])
case("resolve_unknown", PGT + ":422-428", CHAIN, [{"op": "resolve", "arg": "co.fanki.Unknown", "set": []}])
case("resolve_isolated", PGT + ":431-438", [node("co.fanki.Isolated")],
     [{"op": "resolve", "arg": "co.fanki.Isolated", "set": []}])

# ---- analysisOrder -------------------------------------------------------------------------
case("order_chain", PGT + ":443-452", CHAIN + [entry("co.fanki.Controller")],
     [{"op": "analysisOrder", "index": {"co.fanki.Controller": 0},
       "before": [["co.fanki.Service", "co.fanki.Repository"]],
       "derived_list": ["co.fanki.Controller", "co.fanki.Service", "co.fanki.Repository"]}])
case("order_multi_entry", PGT + ":455-476",
     [node("co.fanki.ControllerA"), node("co.fanki.ControllerB"), node("co.fanki.ServiceA"),
      node("co.fanki.ServiceB"), dep("co.fanki.ControllerA", "co.fanki.ServiceA"),
      dep("co.fanki.ControllerB", "co.fanki.ServiceB"), entry("co.fanki.ControllerA"),
      entry("co.fanki.ControllerB")],
     [{"op": "analysisOrder", "size": 4,
       "before": [["co.fanki.ControllerA", "co.fanki.ServiceA"], ["co.fanki.ControllerB", "co.fanki.ServiceB"]],
       "derived_list": ["co.fanki.ControllerA", "co.fanki.ControllerB", "co.fanki.ServiceA", "co.fanki.ServiceB"]}])
case("order_orphans_last", PGT + ":479-491",
     [node("co.fanki.Controller"), node("co.fanki.Service"), node("co.fanki.Orphan"),
      dep("co.fanki.Controller", "co.fanki.Service"), entry("co.fanki.Controller")],
     [{"op": "analysisOrder", "size": 3, "last": "co.fanki.Orphan",
       "derived_list": ["co.fanki.Controller", "co.fanki.Service", "co.fanki.Orphan"]}])
case("order_no_entry_points", PGT + ":493-503", [node("co.fanki.A"), node("co.fanki.B")],
     [{"op": "analysisOrder", "size": 2, "contains": ["co.fanki.A", "co.fanki.B"],
       "derived_list": ["co.fanki.A", "co.fanki.B"]}])
case("order_empty_graph", PGT + ":506-512", [], [{"op": "analysisOrder", "size": 0, "list": []}])
case("order_diamond", PGT + ":515-534",
     [node("co.fanki.Root"), node("co.fanki.Left"), node("co.fanki.Right"), node("co.fanki.Leaf"),
      dep("co.fanki.Root", "co.fanki.Left"), dep("co.fanki.Root", "co.fanki.Right"),
      dep("co.fanki.Left", "co.fanki.Leaf"), dep("co.fanki.Right", "co.fanki.Leaf"), entry("co.fanki.Root")],
     [{"op": "analysisOrder", "size": 4, "distinct": True, "index": {"co.fanki.Root": 0},
       "derived_list": ["co.fanki.Root", "co.fanki.Left", "co.fanki.Right", "co.fanki.Leaf"]}])

# ---- cycles / self loops -------------------------------------------------------------------
CYCLE = [node("co.fanki.A"), node("co.fanki.B"), dep("co.fanki.A", "co.fanki.B"), dep("co.fanki.B", "co.fanki.A")]
case("resolve_cycle", PGT + ":777-790", CYCLE,
     [{"op": "resolve", "arg": "co.fanki.A", "set": ["co.fanki.B"]},
      {"op": "resolve", "arg": "co.fanki.B", "set": ["co.fanki.A"]}])
case("order_cycle", PGT + ":793-807", CYCLE + [entry("co.fanki.A")],
     [{"op": "analysisOrder", "size": 2, "list": ["co.fanki.A", "co.fanki.B"]}])
case("json_roundtrip_cycle", PGT + ":810-826", CYCLE + [entry("co.fanki.A")],
     [{"op": "jsonRoundTrip", "same": ["resolve:co.fanki.A", "resolve:co.fanki.B", "analysisOrder"]}])
case("resolve_self_loop", PGT + ":829-837", [node("co.fanki.A"), dep("co.fanki.A", "co.fanki.A")],
     [{"op": "resolve", "arg": "co.fanki.A", "set": ["co.fanki.A"]},
      {"op": "dependencies", "arg": "co.fanki.A", "derived_set": ["co.fanki.A"]},
      {"op": "dependents", "arg": "co.fanki.A", "derived_set": ["co.fanki.A"]}])

# ---- CodeContextService: stack-trace neighbours (CCS:1550-1578) -----------------------------
case("stacktrace_neighbors", CCT + ":365-456",
     [node("co.fanki.order.OrderController", "OrderController.java"),
      node("co.fanki.order.OrderService", "OrderService.java"),
      dep("co.fanki.order.OrderController", "co.fanki.order.OrderService")],
     [{"op": "resolveNeighbors", "frames": ["co.fanki.order.OrderController"],
       "set": ["co.fanki.order.OrderService"], "first": "co.fanki.order.OrderService"}])
case("stacktrace_all_missing", CCT + ":459-480", [],
     [{"op": "resolveNeighbors", "frames": [], "set": [],
       "note": "frame com.unknown.FooService is not found by the repository, so M is empty and no graph is consulted"}])

# ---- CodeContextService: class dependencies (CCS:1163-1198) ---------------------------------
case("class_dependencies", CCT + ":483-572",
     [node("co.fanki.order.OrderService", "OrderService.java"),
      node("co.fanki.order.OrderRepository", "OrderRepository.java"),
      node("co.fanki.order.OrderController", "OrderController.java"),
      node("co.fanki.order.OrderDto", "OrderDto.java"),
      dep("co.fanki.order.OrderService", "co.fanki.order.OrderRepository"),
      dep("co.fanki.order.OrderController", "co.fanki.order.OrderService"),
      entry("co.fanki.order.OrderController")],
     [{"op": "classDependencies", "arg": "co.fanki.order.OrderService",
       "dependencies": ["co.fanki.order.OrderRepository"],
       "dependents": ["co.fanki.order.OrderController"], "entryPoint": False}])

# ---- CodeContextService: search_project (CCS:1492-1530) -------------------------------------
SEARCH3 = [node("com.example.OrderService", "OrderService.java"),
           node("com.example.OrderController", "OrderController.java"),
           node("com.example.UserService", "UserService.java"),
           entry("com.example.OrderController")]
case("search_matches", CCT + ":823-866", SEARCH3,
     [{"op": "searchProject", "query": "Order", "count": 2, "total": 3,
       "derived_list": ["com.example.OrderService", "com.example.OrderController"]}])
case("search_no_match", CCT + ":887-906", [node("com.example.OrderService", "OrderService.java")],
     [{"op": "searchProject", "query": "ZzzNotExist", "count": 0, "total": 1, "list": []}])

# ---- GraphQueryService (GQS) ----------------------------------------------------------------
case("gq_classes", GQT + ":163-175", GQ_GRAPH, [{"op": "classes", "count": 4}])
case("gq_classes_dependencies", GQT + ":178-199", GQ_GRAPH,
     [{"op": "classesDependencies", "arg": "co.fanki.UserController", "contains": ["co.fanki.UserService"]}])
case("gq_entrypoints", GQT + ":222-236", GQ_GRAPH,
     [{"op": "entrypoints", "count": 1, "list": ["co.fanki.UserController"]}])
case("gq_vertex_simple_name", GQT + ":258-271", GQ_GRAPH,
     [{"op": "resolveClassName", "arg": "UserService", "eq": "co.fanki.UserService"}])
case("gq_vertex_full_name", GQT + ":274-285", GQ_GRAPH,
     [{"op": "resolveClassName", "arg": "co.fanki.UserService", "eq": "co.fanki.UserService"}])
case("gq_vertex_unknown", GQT + ":288-298", GQ_GRAPH,
     [{"op": "resolveClassName", "arg": "Unknown", "eq": None}])
case("gq_vertex_dependencies", GQT + ":329-343", GQ_GRAPH,
     [{"op": "vertexDeps", "arg": "UserController", "outgoing": True, "count": 1,
       "list": ["co.fanki.UserService"]}])
case("gq_vertex_dependents", GQT + ":345-358", GQ_GRAPH,
     [{"op": "vertexDeps", "arg": "UserService", "outgoing": False, "count": 1,
       "list": ["co.fanki.UserController"]}])
case("gq_check_exists", GQT + ":414-428", GQ_GRAPH,
     [{"op": "methodCheck", "arg": "UserController", "check": "createUser", "exists": True,
       "methodName": "createUser"}], methods=GQ_METHODS)
case("gq_check_missing", GQT + ":431-443", GQ_GRAPH,
     [{"op": "methodCheck", "arg": "UserController", "check": "deleteUser", "exists": False}],
     methods=GQ_METHODS)

# ---- graph_query through the MCP tool (GraphQueryMcpToolTest) -------------------------------
case("mcp_classes", MCT + ":135-148", MCP_GRAPH,
     [{"op": "classes", "count": 3,
       "derived_list": ["co.fanki.OrderController", "co.fanki.OrderService", "co.fanki.OrderRepository"]}])
case("mcp_methods_navigation_resolves_simple_name", MCT + ":150-163", MCP_GRAPH,
     [{"op": "resolveClassName", "arg": "OrderController", "eq": "co.fanki.OrderController"}])

# ---- JavaSourceParser-built graphs (graph shape only; the parser itself stays in the host) --
case("parser_edges", JPT + ":98-125",
     [node("co.fanki.app.OrderRepository", "src/main/java/co/fanki/app/OrderRepository.java"),
      node("co.fanki.app.OrderService", "src/main/java/co/fanki/app/OrderService.java"),
      dep("co.fanki.app.OrderService", "co.fanki.app.OrderRepository")],
     [{"op": "resolve", "arg": "co.fanki.app.OrderService", "contains": ["co.fanki.app.OrderRepository"]}])
case("parser_filters_external_imports", JPT + ":206-235",
     [node("co.fanki.app.MyRepository", "src/main/java/co/fanki/app/MyRepository.java"),
      node("co.fanki.app.MyService", "src/main/java/co/fanki/app/MyService.java"),
      dep("co.fanki.app.MyService", "java.util.List"),
      dep("co.fanki.app.MyService", "org.springframework.stereotype.Service"),
      dep("co.fanki.app.MyService", "co.fanki.app.MyRepository")],
     [{"op": "resolve", "arg": "co.fanki.app.MyService", "size": 1, "contains": ["co.fanki.app.MyRepository"]}])
case("parser_static_import", JPT + ":238-265",
     [node("co.fanki.app.Constants", "src/main/java/co/fanki/app/Constants.java"),
      node("co.fanki.app.Printer", "src/main/java/co/fanki/app/Printer.java"),
      dep("co.fanki.app.Printer", "co.fanki.app.Constants")],
     [{"op": "resolve", "arg": "co.fanki.app.Printer", "contains": ["co.fanki.app.Constants"]}])


def main():
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_tests.json")
    with open(out, "w") as f:
        json.dump({"about": "transcribed from waabox/domain-mcp-server unit tests; see transcribe_reference_tests.py",
                   "cases": CASES}, f, indent=1)
    print("wrote %s: %d cases" % (out, len(CASES)))


if __name__ == "__main__":
    main()
