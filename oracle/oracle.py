"""TEST INFRASTRUCTURE -- ctypes binding of oracle/liboracle.so (see oracle.cc's header).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this.  Never imported by the dmgraph package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.ora_ref_new.restype = C.c_void_p
        L.ora_ref_free.argtypes = [C.c_void_p]
        for f in ("ora_ref_add_node", "ora_ref_add_dependency"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        L.ora_ref_mark_entry.argtypes = [C.c_void_p, C.c_char_p]
        L.ora_ref_node_count.argtypes = [C.c_void_p]
        L.ora_ref_node_count.restype = C.c_uint32
        L.ora_ref_entry_count.argtypes = [C.c_void_p]
        L.ora_ref_entry_count.restype = C.c_uint32
        L.ora_ref_contains.argtypes = [C.c_void_p, C.c_char_p]
        L.ora_ref_is_entry.argtypes = [C.c_void_p, C.c_char_p]
        L.ora_ref_source_file.argtypes = [C.c_void_p, C.c_char_p]
        L.ora_ref_source_file.restype = C.c_char_p
        L.ora_ref_name.argtypes = [C.c_void_p, C.c_uint32]
        L.ora_ref_name.restype = C.c_char_p
        L.ora_ref_query.argtypes = [C.c_void_p, C.c_int, C.c_char_p, u32p, C.c_uint32]
        L.ora_ref_query.restype = C.c_uint32
        L.ora_ref_resolve_class_name.argtypes = [C.c_void_p, C.c_char_p]
        L.ora_ref_resolve_class_name.restype = C.c_int64
        L.ora_ref_khop.argtypes = [C.c_void_p, u32p, C.c_uint32, C.c_int, C.c_int, C.c_int, u32p, u8p, C.c_uint32]
        L.ora_ref_khop.restype = C.c_uint32
        L.ora_refsvc_new.argtypes = [C.c_uint32, u64p, u8p, C.c_uint64, u32p, u32p, C.c_uint32, u32p,
                                     C.c_uint32, u32p]
        L.ora_refsvc_new.restype = C.c_void_p
        L.ora_refsvc_free.argtypes = [C.c_void_p]
        L.ora_refsvc_khop_batch.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint64,
                                            u32p, u32p, C.c_int, C.POINTER(C.c_uint64)]
        L.ora_refsvc_khop_batch.restype = C.c_uint64
        L.ora_refsvc_search_batch.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint64, u32p, u8p, C.c_int]
        L.ora_refsvc_search_batch.restype = C.c_uint64
        L.ora_csr_new.argtypes = [C.c_uint32, C.c_uint64, u32p, u32p, C.c_uint32, u32p, C.c_uint32, u32p]
        L.ora_csr_new.restype = C.c_void_p
        L.ora_csr_free.argtypes = [C.c_void_p]
        for f in ("ora_csr_edges", "ora_csr_dropped", "ora_csr_dups"):
            getattr(L, f).argtypes = [C.c_void_p]
            getattr(L, f).restype = C.c_uint64
        L.ora_csr_adjacency_batch.argtypes = [C.c_void_p, C.c_int, C.c_uint64, u32p, u64p, C.c_void_p, C.c_uint64]
        L.ora_csr_adjacency_batch.restype = C.c_uint64
        L.ora_csr_khop_batch.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint64, u32p, u32p, u64p,
                                         C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_uint64)]
        L.ora_csr_khop_batch.restype = C.c_uint64
        L.ora_csr_closure.argtypes = [C.c_void_p, C.c_int, C.c_uint32, u32p, u32p, C.POINTER(C.c_uint64)]
        L.ora_csr_closure.restype = C.c_uint64
        L.ora_csr_analysis_order.argtypes = [C.c_void_p, u32p]
        L.ora_csr_entry_points.argtypes = [C.c_void_p, u32p, C.c_uint32]
        L.ora_csr_entry_points.restype = C.c_uint32
        L.ora_search_batch.argtypes = [C.c_uint32, C.c_uint32, u64p, u8p, C.c_int, C.c_uint64, u32p, u8p, u64p,
                                       C.c_void_p, C.c_uint64, C.c_int]
        L.ora_search_batch.restype = C.c_uint64
        L.ora_search_batch_ac.argtypes = [C.c_uint32, C.c_uint32, u64p, u8p, C.c_uint64, u32p, u8p, u64p,
                                          C.c_void_p, C.c_uint64, C.c_int]
        L.ora_search_batch_ac.restype = C.c_uint64
        L.ora_hardware_threads.restype = C.c_int
        _LIB = L
    return _LIB


def _b(s):
    return None if s is None else s.encode("utf-8")


def _u32(a):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    return a if a.size else np.zeros(1, np.uint32)[:0].copy()


def hardware_threads():
    return lib().ora_hardware_threads()


class RefGraph:
    """String-keyed restatement of ProjectGraph (+ the CCS/GQS scans); names in, names out."""

    def __init__(self):
        self.L = lib()
        self.h = self.L.ora_ref_new()
        self.methods = {}

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ora_ref_free(self.h)
            self.h = None

    def add_node(self, i, f):
        if self.L.ora_ref_add_node(self.h, _b(i), _b(f)):
            raise ValueError("IllegalArgumentException")

    def add_dependency(self, a, b):
        if self.L.ora_ref_add_dependency(self.h, _b(a), _b(b)):
            raise ValueError("IllegalArgumentException")

    def mark_as_entry_point(self, i):
        if self.L.ora_ref_mark_entry(self.h, _b(i)):
            raise ValueError("IllegalArgumentException")

    def node_count(self):
        return self.L.ora_ref_node_count(self.h)

    def entry_point_count(self):
        return self.L.ora_ref_entry_count(self.h)

    def contains(self, i):
        return bool(self.L.ora_ref_contains(self.h, _b(i)))

    def is_entry_point(self, i):
        return bool(self.L.ora_ref_is_entry(self.h, _b(i)))

    def source_file(self, i):
        r = self.L.ora_ref_source_file(self.h, _b(i))
        return None if r is None else r.decode()

    def _names(self, op, arg=None):
        cap = max(1, self.node_count())
        out = np.zeros(cap, np.uint32)
        n = self.L.ora_ref_query(self.h, op, _b(arg), out, cap)
        return [self.L.ora_ref_name(self.h, int(p)).decode() for p in out[:n]]

    def dependencies(self, i):
        return self._names(0, i)

    def dependents(self, i):
        return self._names(1, i)

    def resolve(self, i):
        return self._names(2, i)

    def analysis_order(self):
        return self._names(3)

    def entry_points(self):
        return self._names(4)

    def identifiers(self):
        return self._names(5)

    def search_project(self, query):
        if query is None:
            raise TypeError("query is null")
        return self._names(6, query), self.node_count()

    def resolve_class_name(self, name):
        p = self.L.ora_ref_resolve_class_name(self.h, _b(name))
        return None if p < 0 else self.L.ora_ref_name(self.h, int(p)).decode()

    def khop(self, sources, k, mode="both", exclude_sources=True):
        ids = {n: i for i, n in enumerate(self.identifiers())}
        src = _u32([ids[s] for s in sources if s in ids])
        cap = max(1, self.node_count())
        oi, oh = np.zeros(cap, np.uint32), np.zeros(cap, np.uint8)
        n = self.L.ora_ref_khop(self.h, src, src.size, k, {"out": 0, "in": 1, "both": 2}[mode],
                                int(exclude_sources), oi, oh, cap)
        return {self.L.ora_ref_name(self.h, int(p)).decode(): int(h) for p, h in zip(oi[:n], oh[:n])}

    def resolve_neighbors(self, frames):
        return list(self.khop(frames, 1, "both", True))


class RefService:
    """Many string-keyed project graphs built from arrays; used for CPU-baseline timing."""

    def __init__(self, name_off, name_bytes, src, dst, entry, part_off):
        self.L = lib()
        V = name_off.size - 1
        self.h = self.L.ora_refsvc_new(V, np.ascontiguousarray(name_off, np.uint64),
                                       np.ascontiguousarray(name_bytes, np.uint8), src.size, _u32(src),
                                       _u32(dst), len(entry), _u32(entry), len(part_off) - 1, _u32(part_off))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ora_refsvc_free(self.h)
            self.h = None

    def khop_batch(self, mode, k, exclude, q_begin, q_end, q_off, q_src, threads):
        chk = C.c_uint64(0)
        n = self.L.ora_refsvc_khop_batch(self.h, mode, k, int(exclude), q_begin, q_end, _u32(q_off),
                                         _u32(q_src), threads, C.byref(chk))
        return n, chk.value

    def search_batch(self, project, q_begin, q_end, q_off, q_bytes, threads):
        return self.L.ora_refsvc_search_batch(self.h, project, q_begin, q_end, _u32(q_off),
                                              np.ascontiguousarray(q_bytes, np.uint8), threads)


class CsrOracle:
    """Integer-CSR restatement: ids in, ids out."""

    def __init__(self, n_vertices, src, dst, entry=(), part_off=None):
        self.L = lib()
        self.V = int(n_vertices)
        po = _u32(part_off if part_off is not None else [0, n_vertices])
        self.h = self.L.ora_csr_new(self.V, len(src), _u32(src), _u32(dst), len(entry), _u32(entry),
                                    po.size - 1, po)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ora_csr_free(self.h)
            self.h = None

    @property
    def n_edges(self):
        return self.L.ora_csr_edges(self.h)

    def adjacency_batch(self, direction, ids):
        ids = _u32(ids)
        off = np.zeros(ids.size + 1, np.uint64)
        need = self.L.ora_csr_adjacency_batch(self.h, direction, ids.size, ids, off, None, 0)
        out = np.zeros(max(1, need), np.uint32)
        self.L.ora_csr_adjacency_batch(self.h, direction, ids.size, ids, off, out.ctypes.data, need)
        return off, out[:need]

    def khop_batch(self, mode, k, exclude, q_off, q_src, threads=1, sizes_only=False):
        q_off, q_src = _u32(q_off), _u32(q_src)
        nq = q_off.size - 1
        off = np.zeros(nq + 1, np.uint64)
        te = C.c_uint64(0)
        need = self.L.ora_csr_khop_batch(self.h, mode, k, int(exclude), nq, q_off, q_src, off, None, None, 0,
                                         threads, C.byref(te))
        if sizes_only:
            return off, None, None, te.value
        ids = np.zeros(max(1, need), np.uint32)
        hop = np.zeros(max(1, need), np.uint8)
        self.L.ora_csr_khop_batch(self.h, mode, k, int(exclude), nq, q_off, q_src, off, ids.ctypes.data,
                                  hop.ctypes.data, need, threads, C.byref(te))
        return off, ids[:need], hop[:need], te.value

    def closure(self, direction, sources):
        s = _u32(sources)
        level = np.zeros(max(1, self.V), np.uint32)
        te = C.c_uint64(0)
        reached = self.L.ora_csr_closure(self.h, direction, s.size, s, level, C.byref(te))
        return level[:self.V], reached, te.value

    def analysis_order(self):
        order = np.zeros(max(1, self.V), np.uint32)
        self.L.ora_csr_analysis_order(self.h, order)
        return order[:self.V]

    def entry_points(self):
        out = np.zeros(max(1, self.V), np.uint32)
        n = self.L.ora_csr_entry_points(self.h, out, out.size)
        return out[:n]


def search_batch(name_off, lower_bytes, queries, kind=0, v_lo=0, v_hi=None, threads=1):
    """queries: list of bytes (pre-lowered). Returns (off u64[nq+1], ids u32[])."""
    L = lib()
    nq = len(queries)
    q_off = np.zeros(nq + 1, np.uint32)
    np.cumsum([len(q) for q in queries], out=q_off[1:])
    qb = np.frombuffer(b"".join(queries) + b"\0", dtype=np.uint8).copy()
    if v_hi is None:
        v_hi = name_off.size - 1
    off = np.zeros(nq + 1, np.uint64)
    lb = np.ascontiguousarray(lower_bytes, np.uint8)
    if lb.size == 0:
        lb = np.zeros(1, np.uint8)
    no = np.ascontiguousarray(name_off, np.uint64)
    need = L.ora_search_batch(v_lo, v_hi, no, lb, kind, nq, q_off, qb, off, None, 0, threads)
    ids = np.zeros(max(1, need), np.uint32)
    L.ora_search_batch(v_lo, v_hi, no, lb, kind, nq, q_off, qb, off, ids.ctypes.data, need, threads)
    return off, ids[:need]


def search_batch_ac(name_off, lower_bytes, queries, v_lo=0, v_hi=None, threads=1):
    """kind-0 search (all hits, nodes order) by one Aho-Corasick pass over the arena: the same rows as
    search_batch(kind=0), fast enough for BASELINE config 3 at full size.  The CPU tests hold it to
    search_batch (the restatement of the reference's per-identifier contains loop)."""
    L = lib()
    nq = len(queries)
    q_off = np.zeros(nq + 1, np.uint32)
    np.cumsum([len(q) for q in queries], out=q_off[1:])
    qb = np.frombuffer(b"".join(queries) + b"\0", dtype=np.uint8).copy()
    if v_hi is None:
        v_hi = name_off.size - 1
    off = np.zeros(nq + 1, np.uint64)
    lb = np.ascontiguousarray(lower_bytes, np.uint8)
    if lb.size == 0:
        lb = np.zeros(1, np.uint8)
    no = np.ascontiguousarray(name_off, np.uint64)
    need = L.ora_search_batch_ac(v_lo, v_hi, no, lb, nq, q_off, qb, off, None, 0, threads)
    ids = np.zeros(max(1, need), np.uint32)
    L.ora_search_batch_ac(v_lo, v_hi, no, lb, nq, q_off, qb, off, ids.ctypes.data, need, threads)
    return off, ids[:need]
