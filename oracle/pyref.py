"""TEST INFRASTRUCTURE -- not product code.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module.

Pure-Python, structure-faithful restatement of the reference's hot path, written
independently of the C++ oracle (oracle/oracle.cc) so each can check the other on small
inputs.  Python dicts keep insertion order, which is what java.util.LinkedHashMap /
LinkedHashSet give the reference; java.util.HashMap iteration order (unpinned by any
reference test) is replaced by "whatever order", and every caller compares
those results as sets or canonically sorted.

Parity status: pinned against tests/golden/reference_tests.json (the reference's own unit
tests, transcribed) by tests/test_oracle_golden.py.  The reference itself (Java 21) cannot
be run in this image: no JVM.

Citations: PG = src/main/java/co/fanki/domainmcp/analysis/domain/ProjectGraph.java,
CCS = .../analysis/application/CodeContextService.java,
GQS = .../analysis/application/GraphQueryService.java,
GQ = .../analysis/application/GraphQuery.java, all under /root/reference.
"""
import json

_JAVA_WS = set(range(9, 14)) | set(range(28, 33))   # Character.isWhitespace, ASCII part


def _is_blank(s):
    """String.isBlank() (shared/Preconditions.java:41-47)."""
    return all((ord(c) in _JAVA_WS) or (ord(c) > 127 and c.isspace()) for c in s)


def _require_non_blank(v, msg):
    if v is None or _is_blank(v):
        raise ValueError(msg)   # IllegalArgumentException


def ascii_lower(s):
    """toLowerCase() restricted to ASCII; non-ASCII lowering is the host's job."""
    return "".join(chr(ord(c) + 32) if "A" <= c <= "Z" else c for c in s)


def extract_simple_name(fqcn):
    """CCS:1538-1541 / GQS:458-461: text after the last '.'."""
    k = fqcn.rfind(".")
    return fqcn[k + 1:] if k >= 0 else fqcn


class ProjectGraph:
    """PG:88-125 -- nodes (LinkedHashMap), edges (HashMap of LinkedHashSet), entryPoints."""

    def __init__(self):
        self.nodes = {}          # id -> sourceFile, insertion ordered
        self.edges = {}          # id -> {target: None}, targets insertion ordered
        self.entry = {}          # insertion-ordered set
        self.methods = {}        # id -> [method names] (MethodInfo.methodName only)

    # -- build (PG:160-197) -----------------------------------------------------------------
    def add_node(self, identifier, source_file):
        _require_non_blank(identifier, "Identifier is required")
        _require_non_blank(source_file, "Source file is required")
        self.nodes[identifier] = source_file      # re-put keeps position, overwrites value

    def add_dependency(self, frm, to):
        _require_non_blank(frm, "From identifier is required")
        _require_non_blank(to, "To identifier is required")
        if frm not in self.nodes or to not in self.nodes:
            return
        self.edges.setdefault(frm, {})[to] = None

    def mark_as_entry_point(self, identifier):
        _require_non_blank(identifier, "Entry point identifier is required")
        if identifier in self.nodes:
            self.entry[identifier] = None

    # -- reads (PG:299-493) -----------------------------------------------------------------
    def dependencies(self, identifier):
        if identifier is None or identifier not in self.nodes:
            return []
        return list(self.edges.get(identifier, {}))

    def dependents(self, identifier):
        if identifier is None or identifier not in self.nodes:
            return []
        return [s for s, t in self.edges.items() if identifier in t]

    def resolve(self, identifier):
        if identifier is None or identifier not in self.nodes:
            return []
        out = dict.fromkeys(self.edges.get(identifier, {}))
        for s, t in self.edges.items():
            if identifier in t:
                out[s] = None
        return list(out)

    def is_entry_point(self, identifier):
        return identifier is not None and identifier in self.entry

    def entry_points(self):
        return list(self.entry)

    def entry_point_count(self):
        return len(self.entry)

    def analysis_order(self):
        """PG:390-421: FIFO BFS over out-edges from the entry points, then the rest in node order."""
        if not self.nodes:
            return []
        visited = {}
        queue = []
        for ep in self.entry:
            if ep not in visited:
                visited[ep] = None
                queue.append(ep)
        head = 0
        while head < len(queue):
            cur = queue[head]
            head += 1
            for d in self.edges.get(cur, {}):
                if d not in visited:
                    visited[d] = None
                    queue.append(d)
        for i in self.nodes:
            visited.setdefault(i, None)
        return list(visited)

    def source_file(self, identifier):
        return self.nodes.get(identifier)

    def contains(self, identifier):
        return identifier in self.nodes

    def node_count(self):
        return len(self.nodes)

    def identifiers(self):
        return list(self.nodes)

    # -- wire format (PG:634-905), graph sections only --------------------------------------
    def to_json(self):
        return json.dumps({
            "nodes": {i: {"sourceFile": f} for i, f in self.nodes.items()},
            "edges": {s: list(t) for s, t in self.edges.items()},
            "entryPoints": list(self.entry),
        })

    @staticmethod
    def from_json(text):
        """PG:758-905. Note: edges are NOT re-validated against nodes (PG:782-795)."""
        root = json.loads(text)
        g = ProjectGraph()
        for i, n in root.get("nodes", {}).items():
            g.nodes[i] = n.get("sourceFile")
        for s, ts in root.get("edges", {}).items():
            g.edges[s] = dict.fromkeys(ts)
        for e in root.get("entryPoints", []):
            g.entry[e] = None
        return g


# -- CodeContextService graph halves -------------------------------------------------------------
def search_project(graph, query):
    """CCS:1492-1530: ordered hits + total. query None -> NPE in Java -> TypeError here."""
    if query is None:
        raise TypeError("query is null")
    lq = ascii_lower(query)
    hits = []
    for ident in graph.identifiers():
        simple = extract_simple_name(ident)
        if lq in ascii_lower(ident) or lq in ascii_lower(simple):
            hits.append(ident)
    return hits, graph.node_count()


def resolve_neighbors(graph, matched_class_names):
    """CCS:1558-1578 (graph part): N = (U resolve(m)) minus M, M distinct."""
    if graph is None:
        return []
    m = dict.fromkeys(matched_class_names)
    n = {}
    for name in m:
        for nb in graph.resolve(name):
            if nb not in m:
                n[nb] = None
    return list(n)


def khop(graph, sources, k, mode="both", exclude_sources=True):
    """SURVEY 8a T3/T4 extension: iterate the reference's own 1-hop primitive level by level.

    L0 = distinct known sources; L(i+1) = (U_{u in Li} step(u)) minus U_{j<=i} Lj.
    Returns {id: hop}.  With exclude_sources=False a source that is reachable from the
    source set in >=1 step is reported at the hop it is first re-reached (self-loop -> 1),
    which at k=1 with one source gives exactly resolve()/dependencies()/dependents().
    """
    step = {"both": graph.resolve, "out": graph.dependencies, "in": graph.dependents}[mode]
    srcs = [s for s in dict.fromkeys(sources) if s is not None and graph.contains(s)]
    seen = set(srcs) if exclude_sources else set()
    hop = {}
    frontier = list(srcs)
    for lvl in range(1, k + 1):
        nxt = []
        for u in frontier:
            for v in step(u):
                if v not in seen:
                    seen.add(v)
                    hop[v] = lvl
                    nxt.append(v)
        frontier = nxt
        if not frontier:
            break
    return hop


def class_dependencies(graph, class_name):
    """CCS:1163-1198 (graph part): ordered dependencies, dependents set, entry-point flag."""
    return graph.dependencies(class_name), graph.dependents(class_name), graph.is_entry_point(class_name)


# -- GraphQueryService graph halves --------------------------------------------------------------
def resolve_class_name(graph, class_name):
    """GQS:423-456."""
    if graph.contains(class_name):
        return class_name
    low = ascii_lower(class_name)
    best = None
    for ident in graph.identifiers():
        if ascii_lower(extract_simple_name(ident)) == low:
            if best is not None:
                return best
            best = ident
    if best is not None:
        return best
    for ident in graph.identifiers():
        if low in ascii_lower(ident):
            return ident
    return None


def _jtrim(s):
    """String.trim(): strips code units <= U+0020 at both ends."""
    a, b = 0, len(s)
    while a < b and s[a] <= " ":
        a += 1
    while b > a and s[b - 1] <= " ":
        b -= 1
    return s[a:b]


class InvalidQuery(ValueError):
    """DomainException(..., "INVALID_QUERY")."""


def parse_graph_query(text):
    """GQ:95-163. Returns (project, [(kind, value)...]) with kind in NAVIGATE/INCLUDE/CHECK."""
    if text is None or _is_blank(text):
        raise InvalidQuery("Query is required")
    parts = _jtrim(text).split(":")
    while parts and parts[-1] == "":      # Java split drops trailing empty strings
        parts.pop()
    if len(parts) < 2:
        raise InvalidQuery("Query must have at least project:target")
    project = _jtrim(parts[0])
    if not project:
        raise InvalidQuery("Project name is required")
    segs = []
    for p in parts[1:]:
        p = _jtrim(p)
        if not p:
            continue                       # GQ:122-124: empty segments are skipped
        if p.startswith("+"):
            v = _jtrim(p[1:])
            if not v:
                raise InvalidQuery("Include modifier (+) requires a value")
            segs.append(("INCLUDE", v))
        elif p.startswith("?"):
            v = _jtrim(p[1:])
            if not v:
                raise InvalidQuery("Check (?) requires a value")
            segs.append(("CHECK", v))
        else:
            segs.append(("NAVIGATE", p))
    if not segs:
        raise InvalidQuery("Query must have at least one target after project")
    if segs[0][0] != "NAVIGATE":
        raise InvalidQuery("First segment after project must be a navigation target")
    return project, segs


def method_check(graph, identifier, check_value):
    """GQS:362-394: case-insensitive equality over the vertex's method names; first match."""
    low = ascii_lower(check_value)
    for m in graph.methods.get(identifier, []):
        if ascii_lower(m) == low:
            return True, m
    return False, None


# ---- CodeContextService.getStackTraceContext, graph half (CCS:1392-1446 + 1550-1578) ---------------
FRAME_MISSING, FRAME_CLASS_MISSING, FRAME_FOUND = 0, 1, 2


def stack_trace_context(graphs, methods, frames, k=1):
    """graphs: list of (project_id, ProjectGraph) in registry order; methods: {class name: set of
    method names} (source_methods rows; matching is SQL equality, case sensitive); frames: list of
    (class_name, method_name).  Returns (statuses, neighbours {name: hop}).
      - a frame whose class no project holds is MISSING (CCS:1405-1410); whose class is known but
        whose method is not, CLASS_MISSING (CCS:1415-1423); else FOUND
      - the project of the FIRST found frame is the graph neighbours are resolved in (CCS:1426-1434)
      - M = class names of found frames; neighbours = union of resolve(m) minus M, on that one graph
        (CCS:1558-1578): a found frame of another project takes part BY NAME only
      - k > 1 iterates the same step level by level; k = 1 is the reference."""
    status, project = [], None
    for cls, meth in frames:
        owner = next((pid for pid, g in graphs if g.contains(cls)), None)
        if owner is None:
            status.append(FRAME_MISSING)
        elif methods is not None and meth not in methods.get(cls, ()):
            status.append(FRAME_CLASS_MISSING)
        else:
            status.append(FRAME_FOUND)
            if project is None:
                project = owner
    if project is None:
        return status, {}
    graph = next(g for pid, g in graphs if pid == project)
    matched = [c for (c, _m), s in zip(frames, status) if s == FRAME_FOUND]
    return status, khop(graph, matched, k, "both", True)
