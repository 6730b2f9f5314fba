"""Names-in / names-out adapter over the C ABI, playing the part of the Java shim: it keeps the
String<->id map (the host owns `nodes`), enforces Preconditions.requireNonBlank, lower-cases
queries, and sends every read to the device through dmgraph (include/dmgraph.h).  It computes
no graph result itself."""
import numpy as np

import dmgraph
from dmgraph import synth

_WS = set(range(9, 14)) | set(range(28, 33))


def _blank(s):
    return s is None or all(ord(c) in _WS for c in s)


class EngineGraph:
    def __init__(self, engine):
        self.engine = engine
        self.names, self.ids, self.files = [], {}, []
        self.src, self.dst, self.entry = [], [], []
        self.methods = {}
        self._g = None

    # -- build: the host-side half of PG:160-197 (the device applies the same rules again) -----
    def add_node(self, i, f):
        if _blank(i) or _blank(f):
            raise ValueError("IllegalArgumentException")
        if i in self.ids:
            self.files[self.ids[i]] = f
        else:
            self.ids[i] = len(self.names)
            self.names.append(i)
            self.files.append(f)
        self._drop()

    def add_dependency(self, a, b):
        if _blank(a) or _blank(b):
            raise ValueError("IllegalArgumentException")
        # unknown endpoints are sent as an out-of-range id: the engine drops the edge (PG:179-181)
        self.src.append(self.ids.get(a, 0xFFFFFFFF))
        self.dst.append(self.ids.get(b, 0xFFFFFFFF))
        self._drop()

    def mark_as_entry_point(self, i):
        if _blank(i):
            raise ValueError("IllegalArgumentException")
        self.entry.append(self.ids.get(i, 0xFFFFFFFF))
        self._drop()

    def _drop(self):
        if self._g is not None:
            self._g.release()
            self._g = None

    def graph(self):
        if self._g is None:
            off, arena = synth.pack_names(self.names)
            self._g = dmgraph.Graph(self.engine, len(self.names), self.src, self.dst, self.entry,
                                    name_off=off, name_bytes=arena)
        return self._g

    def _id(self, name):
        return 0xFFFFFFFF if name is None else self.ids.get(name, 0xFFFFFFFF)

    # -- reads ------------------------------------------------------------------------------
    def node_count(self):
        return self.graph().info["n_vertices"]

    def contains(self, i):
        return i in self.ids

    def identifiers(self):
        return list(self.names)

    def source_file(self, i):
        return self.files[self.ids[i]] if i in self.ids else None

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
        if name in self.ids:
            return name
        q = name.lower().encode("utf-8")
        for kind in (dmgraph.SEARCH_SIMPLE_EQ_FIRST, dmgraph.SEARCH_CONTAINS_FIRST):
            _off, ids = self.graph().search_batch(kind, [q])
            if ids.size:
                return self.names[ids[0]]
        return None
