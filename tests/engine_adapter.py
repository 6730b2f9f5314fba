"""Names-in / names-out adapter over the C ABI, playing the part of the Java shim: it keeps the
String<->id map (the host owns `nodes`), enforces Preconditions.requireNonBlank, lower-cases
queries, and sends every read to the device through dmgraph (include/dmgraph.h).  It computes
no graph result itself."""
import dmgraph

from dmgraph.host import ProjectGraph


class EngineGraph:
    def __init__(self, engine):
        self.engine = engine
        self.pg = ProjectGraph()          # C++ host mirror: String<->id map, build rules (PG:160-197)
        self.methods = {}
        self._g = None

    def add_node(self, i, f):
        self.pg.add_node(i, f)
        self._drop()

    def add_dependency(self, a, b):
        self.pg.add_dependency(a, b)
        self._drop()

    def mark_as_entry_point(self, i):
        self.pg.mark_as_entry_point(i)
        self._drop()

    def _drop(self):
        if self._g is not None:
            self._g.release()
            self._g = None

    def graph(self):
        if self._g is None:
            self._g = self.pg.publish(self.engine)
        return self._g

    @property
    def names(self):
        return _Names(self.pg)

    def _id(self, name):
        v = self.pg.id_of(name)
        return 0xFFFFFFFF if v < 0 else v

    # -- reads ------------------------------------------------------------------------------
    def node_count(self):
        return self.graph().info["n_vertices"]

    def contains(self, i):
        return self.pg.contains(i)

    def identifiers(self):
        return self.pg.identifiers()

    def source_file(self, i):
        return self.pg.source_file(i)

    def entry_points(self):
        return [self.names[v] for v in self.graph().entry_points()]

    def entry_point_count(self):
        return self.graph().info["n_entry"]

    def is_entry_point(self, i):
        return bool(self.graph().is_entry_point_batch([self._id(i)])[0])

    def dependencies(self, i):
        _off, out = self.graph().adjacency_batch(dmgraph.OUT, [self._id(i)])
        return [self.names[v] for v in out]

    def dependents(self, i):
        _off, out = self.graph().adjacency_batch(dmgraph.IN, [self._id(i)])
        return [self.names[v] for v in out]

    def resolve(self, i):
        return list(self.khop([i], 1, "both", False))

    def khop(self, sources, k, mode="both", exclude_sources=True):
        s = [self._id(x) for x in sources]
        _off, ids, hop = self.graph().neighbors_batch(dmgraph.MODES[mode], k, exclude_sources, [0, len(s)], s)
        return {self.names[v]: int(h) for v, h in zip(ids, hop)}

    def resolve_neighbors(self, frames):
        return list(self.khop(frames, 1, "both", True))

    def analysis_order(self):
        return [self.names[v] for v in self.graph().analysis_order()]

    def search_project(self, query):
        if query is None:
            raise TypeError("query is null")
        q = query.lower().encode("utf-8")          # the host's toLowerCase (CCS:1507)
        _off, ids = self.graph().search_batch(dmgraph.SEARCH_CONTAINS_ALL, [q])
        return [self.names[v] for v in ids], self.node_count()

    def resolve_class_name(self, name):
        """GQS:423-456: exact id on the host map, then the two ordered device scans."""
        if self.pg.contains(name):
            return name
        q = name.lower().encode("utf-8")
        for kind in (dmgraph.SEARCH_SIMPLE_EQ_FIRST, dmgraph.SEARCH_CONTAINS_FIRST):
            _off, ids = self.graph().search_batch(kind, [q])
            if ids.size:
                return self.names[ids[0]]
        return None


class _Names:
    def __init__(self, pg):
        self.pg = pg

    def __getitem__(self, v):
        return self.pg.name_of(int(v))


class EngineQueryService:
    """GraphQueryService.execute (GQS:53-73) over dmg_query_batch: the lexer and the exact-id lookup
    stay on the host (here: oracle-free Python), every graph read is one batched device call."""

    def __init__(self, eg, methods=None):
        self.eg = eg
        self.methods = methods or {}
        names = eg.identifiers()
        eg.graph().set_methods([self.methods.get(n, []) for n in names])
        self.catalog = [m for n in names for m in self.methods.get(n, [])]

    def resolve(self, target):
        if self.eg.pg.contains(target):
            return self.eg.pg.id_of(target)
        (ids, aux), = self.eg.graph().query_batch([{"op": dmgraph.OP_RESOLVE_NAME, "text": target.lower().encode()}])
        return ids[0] if ids else None

    def execute(self, tokens):
        """tokens: [(kind, value)] after the project segment, as GraphQuery.parse yields them."""
        g, names = self.eg.graph(), self.eg.names
        navs = [v for k, v in tokens if k == "NAVIGATE"]
        checks = [v for k, v in tokens if k == "CHECK"]
        head = navs[0].lower()
        if head == "entrypoints":
            (ids, _), = g.query_batch([{"op": dmgraph.OP_ENTRYPOINTS}])
            return "entrypoints", [names[i] for i in ids]
        if head == "classes":
            return "classes", self.eg.identifiers()
        v = self.resolve(navs[0])
        if v is None:
            raise LookupError("CLASS_NOT_FOUND")
        if checks:
            (ids, _), = g.query_batch([{"op": dmgraph.OP_METHOD, "vertex": v, "text": checks[0].lower().encode()}])
            return "check", (bool(ids), self.catalog[ids[0]] if ids else None)
        sub = navs[1].lower() if len(navs) > 1 else None
        if sub == "dependencies":
            (ids, _), = g.query_batch([{"op": dmgraph.OP_DEPENDENCIES, "vertex": v}])
            return "dependencies", [names[i] for i in ids]
        if sub == "dependents":
            (ids, _), = g.query_batch([{"op": dmgraph.OP_DEPENDENTS, "vertex": v}])
            return "dependents", [names[i] for i in ids]
        if sub == "method":
            if len(navs) < 3:
                raise ValueError("INVALID_QUERY")
            (ids, _), = g.query_batch([{"op": dmgraph.OP_METHOD, "vertex": v, "text": navs[2].lower().encode()}])
            if not ids:
                raise LookupError("METHOD_NOT_FOUND")
            return "method", self.catalog[ids[0]]
        if sub == "methods":
            return "methods", self.methods.get(names[v], [])
        (ids, aux), = g.query_batch([{"op": dmgraph.OP_OVERVIEW, "vertex": v}])
        return "class", {"className": names[v], "dependencies": [names[i] for i, a in zip(ids, aux) if a == 0],
                         "dependents": [names[i] for i, a in zip(ids, aux) if a == 1]}
