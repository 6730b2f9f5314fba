"""ctypes binding of include/dmgraph_host.h: the C++ host-side mirror of ProjectGraph (build half,
wire format) and GraphService.  Reads that compute a graph result go to the device through
dmgraph.Graph; this module only moves strings and ids."""
import ctypes as C

import dmgraph
from dmgraph import DmgError, Graph, _check, lib


def _bind():
    L = lib()
    if getattr(L, "_host_bound", False):
        return L
    vp, cp, i32, u32, u64, i64 = C.c_void_p, C.c_char_p, C.c_int32, C.c_uint32, C.c_uint64, C.c_int64
    sig = {
        "dmg_pg_new": (vp, []), "dmg_pg_free": (None, [vp]),
        "dmg_pg_add_node": (i32, [vp, cp, cp]), "dmg_pg_add_dependency": (i32, [vp, cp, cp]),
        "dmg_pg_mark_entry_point": (i32, [vp, cp]), "dmg_pg_node_count": (u32, [vp]),
        "dmg_pg_entry_point_count": (u32, [vp]), "dmg_pg_edge_count": (u64, [vp]),
        "dmg_pg_contains": (i32, [vp, cp]), "dmg_pg_id_of": (i64, [vp, cp]),
        "dmg_pg_name_of": (cp, [vp, u32]), "dmg_pg_source_file": (cp, [vp, cp]),
        "dmg_pg_publish": (i32, [vp, vp, C.POINTER(vp)]),
        "dmg_pg_to_json": (i32, [vp, C.POINTER(vp)]), "dmg_pg_from_json": (i32, [cp, C.POINTER(vp)]),
        "dmg_free_string": (None, [vp]),
        "dmg_service_new": (i32, [vp, C.POINTER(vp)]), "dmg_service_free": (None, [vp]),
        "dmg_service_put": (i32, [vp, cp, cp, vp]), "dmg_service_load_json": (i32, [vp, cp, cp, cp]),
        "dmg_service_get": (vp, [vp, cp]), "dmg_service_get_by_name": (vp, [vp, cp]),
        "dmg_service_size": (u32, [vp]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype, f.argtypes = res, args
    L._host_bound = True
    return L


def _b(s):
    return None if s is None else s.encode("utf-8")


def _iae(rc):
    if rc == dmgraph.EINVAL:
        raise ValueError("IllegalArgumentException: " + lib().dmg_last_error().decode())
    _check(rc)


class _GraphView(Graph):
    """dmgraph.Graph over an existing dmg_graph handle."""

    def __init__(self, engine, handle):
        self.eng, self.L, self.h = engine, engine.L, C.c_void_p(handle)
        info = dmgraph.GraphInfo()
        _check(self.L.dmg_graph_info_get(self.h, C.byref(info)))
        self.info = {f: getattr(info, f) for f, _ in dmgraph.GraphInfo._fields_}
        self.V = self.info["n_vertices"]


class ProjectGraph:
    """Host half of the reference's ProjectGraph (PG:160-197, 454-493, 634-905)."""

    def __init__(self, handle=None):
        self.L = _bind()
        self.h = handle if handle is not None else self.L.dmg_pg_new()

    def __del__(self):
        if getattr(self, "h", None):
            self.L.dmg_pg_free(self.h)
            self.h = None

    def add_node(self, identifier, source_file):
        _iae(self.L.dmg_pg_add_node(self.h, _b(identifier), _b(source_file)))

    def add_dependency(self, frm, to):
        _iae(self.L.dmg_pg_add_dependency(self.h, _b(frm), _b(to)))

    def mark_as_entry_point(self, identifier):
        _iae(self.L.dmg_pg_mark_entry_point(self.h, _b(identifier)))

    def node_count(self):
        return self.L.dmg_pg_node_count(self.h)

    def entry_point_count(self):
        return self.L.dmg_pg_entry_point_count(self.h)

    def edge_count(self):
        return self.L.dmg_pg_edge_count(self.h)

    def contains(self, identifier):
        return identifier is not None and bool(self.L.dmg_pg_contains(self.h, _b(identifier)))

    def id_of(self, identifier):
        return -1 if identifier is None else self.L.dmg_pg_id_of(self.h, _b(identifier))

    def name_of(self, vertex):
        r = self.L.dmg_pg_name_of(self.h, int(vertex))
        return None if r is None else r.decode("utf-8")

    def identifiers(self):
        return [self.name_of(v) for v in range(self.node_count())]

    def source_file(self, identifier):
        r = None if identifier is None else self.L.dmg_pg_source_file(self.h, _b(identifier))
        return None if r is None else r.decode("utf-8")

    def to_json(self):
        p = C.c_void_p()
        _check(self.L.dmg_pg_to_json(self.h, C.byref(p)))
        try:
            return C.string_at(p.value).decode("utf-8")
        finally:
            self.L.dmg_free_string(p)

    @staticmethod
    def from_json(text):
        L = _bind()
        h = C.c_void_p()
        _iae(L.dmg_pg_from_json(_b(text), C.byref(h)))
        return ProjectGraph(h.value)

    def publish(self, engine):
        g = C.c_void_p()
        _check(self.L.dmg_pg_publish(engine.h, self.h, C.byref(g)))
        return _GraphView(engine, g.value)


class GraphService:
    """GraphService (GS:34-146): projectId -> device graph, project-name index, swap on put."""

    def __init__(self, engine):
        self.L, self.engine = _bind(), engine
        self.h = C.c_void_p()
        _check(self.L.dmg_service_new(engine.h, C.byref(self.h)))

    def close(self):
        if self.h:
            self.L.dmg_service_free(self.h)
            self.h = C.c_void_p()

    def put(self, project_id, project_name, pg):
        _check(self.L.dmg_service_put(self.h, _b(project_id), _b(project_name), pg.h))

    def load_json(self, project_id, project_name, text):
        """Returns False when the JSON does not parse (the reference logs and skips, GS:62-72)."""
        rc = self.L.dmg_service_load_json(self.h, _b(project_id), _b(project_name), _b(text))
        if rc == dmgraph.EINVAL:
            return False
        _check(rc)
        return True

    def get_graph(self, project_id):
        h = self.L.dmg_service_get(self.h, _b(project_id))
        return None if not h else _GraphView(self.engine, h)

    def get_graph_by_project_name(self, name):
        h = self.L.dmg_service_get_by_name(self.h, _b(name))
        return None if not h else _GraphView(self.engine, h)

    def size(self):
        return self.L.dmg_service_size(self.h)
