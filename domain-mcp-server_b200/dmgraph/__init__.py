"""ctypes binding of libdmgraph.so (include/dmgraph.h) -- the harness-side stand-in for the
Java shim that INTEGRATION.md describes.  Thin on purpose: every call goes straight to the
C ABI; nothing here computes a result, and there is no CPU fallback -- if the CUDA library is
missing or no device is present the calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DMG_LIB") or os.path.join(os.path.dirname(_HERE), "libdmgraph.so")   # DMG_LIB: tuning builds

OK, EINVAL, ENOMEM, ECAP, ECUDA, ENCCL, ENODEV = range(7)
OUT, IN, BOTH = 0, 1, 2
SEARCH_CONTAINS_ALL, SEARCH_SIMPLE_EQ_FIRST, SEARCH_CONTAINS_FIRST = 0, 1, 2
LEVEL_UNREACHED = 0xFFFFFFFF
LOOKUP_EXACT, LOOKUP_SIMPLE_FIRST = 0, 1
MODES = {"out": OUT, "in": IN, "both": BOTH}


class DmgError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("dmgraph error %d: %s" % (code, msg))
        self.code = code


class EngineOpts(C.Structure):
    _fields_ = [("device", C.c_int32), ("flags", C.c_int32), ("scratch_bytes", C.c_uint64)]


class GraphDesc(C.Structure):
    _fields_ = [("n_vertices", C.c_uint32), ("n_entry", C.c_uint32), ("n_parts", C.c_uint32),
                ("reserved", C.c_uint32), ("n_edges", C.c_uint64),
                ("name_off", C.c_void_p), ("name_bytes", C.c_void_p), ("lower_bytes", C.c_void_p),
                ("edge_src", C.c_void_p), ("edge_dst", C.c_void_p), ("entry_ids", C.c_void_p),
                ("part_off", C.c_void_p)]


class GraphInfo(C.Structure):
    _fields_ = [("n_vertices", C.c_uint32), ("n_entry", C.c_uint32), ("n_parts", C.c_uint32),
                ("max_part_vertices", C.c_uint32), ("n_edges", C.c_uint64), ("dropped_edges", C.c_uint64),
                ("duplicate_edges", C.c_uint64), ("name_bytes", C.c_uint64), ("device_bytes", C.c_uint64)]


class QueryOp(C.Structure):
    _fields_ = [("opcode", C.c_uint32), ("vertex", C.c_uint32), ("mode", C.c_uint32), ("k", C.c_uint32),
                ("exclude_sources", C.c_uint32), ("text_begin", C.c_uint32), ("text_end", C.c_uint32),
                ("project", C.c_uint32)]


OP_DEPENDENCIES, OP_DEPENDENTS, OP_OVERVIEW, OP_KHOP, OP_ENTRYPOINTS, OP_RESOLVE_NAME, OP_METHOD, OP_ENDPOINTS = range(8)


class CallStats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("edges_traversed", C.c_uint64), ("edges_read", C.c_uint64),
                ("rows_expanded", C.c_uint64),
                ("results", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("device_ms", C.c_float), ("reserved", C.c_float)]


class DistClosureStats(C.Structure):
    _fields_ = [("levels", C.c_uint32), ("push_levels", C.c_uint32), ("pull_levels", C.c_uint32),
                ("transport", C.c_uint32), ("edges_read", C.c_uint64), ("exchanged_bytes_per_rank", C.c_uint64),
                ("device_ms", C.c_float), ("reserved2", C.c_float)]


_lib = None


def lib():
    """Loads libdmgraph.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libdmgraph.so not built: run `make -C domain-mcp-server_b200` "
                              "(or __graft_entry__.build())")
        L = C.CDLL(LIB_PATH)
        vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int32
        pu64 = C.POINTER(C.c_uint64)
        sig = {
            "dmg_device_count": (i32, []),
            "dmg_engine_create": (i32, [C.POINTER(EngineOpts), C.POINTER(vp)]),
            "dmg_engine_destroy": (None, [vp]),
            "dmg_last_error": (C.c_char_p, []),
            "dmg_version": (C.c_char_p, []),
            "dmg_graph_create": (i32, [vp, C.POINTER(GraphDesc), C.POINTER(vp)]),
            "dmg_graph_retain": (None, [vp]),
            "dmg_graph_release": (None, [vp]),
            "dmg_graph_save": (i32, [vp, C.c_char_p]),
            "dmg_graph_load": (i32, [vp, C.c_char_p, C.POINTER(vp)]),
            "dmg_graph_info_get": (i32, [vp, C.POINTER(GraphInfo)]),
            "dmg_adjacency_batch": (i32, [vp, i32, u64, vp, vp, vp, u64, pu64]),
            "dmg_neighbors_batch": (i32, [vp, i32, i32, i32, u64, vp, vp, vp, vp, vp, u64, pu64]),
            "dmg_neighbors_batch_dev": (i32, [vp, i32, i32, i32, u64, vp, vp, vp, vp, vp, u64, pu64]),
            "dmg_neighbors_batch_rows": (i32, [vp, i32, i32, i32, u64, vp, vp, C.POINTER(u32), vp, vp, u64, pu64]),
            "dmg_neighbors_batch_dev_static": (i32, [vp, i32, i32, i32, u64, vp, vp, vp, vp, vp, u64, pu64]),
            "dmg_neighbors_batch_bits": (i32, [vp, i32, i32, i32, u64, vp, vp, C.POINTER(u32), C.POINTER(u32), vp, vp,
                                                vp, u64, pu64]),
            "dmg_stacktrace_batch": (i32, [vp, i32, u64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, u64, pu64]),
            "dmg_graph_set_method_names": (i32, [vp, vp, vp]),
            "dmg_dist_create": (i32, [vp, i32, C.POINTER(vp)]),
            "dmg_dist_destroy": (None, [vp]),
            "dmg_dist_size": (i32, [vp]),
            "dmg_dist_engine": (i32, [vp, i32, C.POINTER(vp)]),
            "dmg_dist_graph_create": (i32, [vp, C.POINTER(GraphDesc), i32, C.POINTER(vp)]),
            "dmg_dist_graph_release": (None, [vp]),
            "dmg_dist_graph_part": (i32, [vp, i32, C.POINTER(vp), C.POINTER(u32), C.POINTER(u32)]),
            "dmg_dist_neighbors_batch": (i32, [vp, i32, i32, i32, u64, vp, vp, vp, vp, vp, u64, pu64]),
            "dmg_dist_closure": (i32, [vp, i32, u32, vp, vp, pu64, C.POINTER(DistClosureStats)]),
            "dmg_dist_analysis_order": (i32, [vp, vp]),
            "dmg_dist_query_batch": (i32, [vp, u64, vp, vp, vp, vp, vp, u64, pu64]),
            "dmg_bits_decode": (i32, [u32, u32, u64, vp, vp, vp, vp, vp, vp, u64, pu64, i32]),
            "dmg_closure": (i32, [vp, i32, u32, vp, vp, pu64]),
            "dmg_analysis_order": (i32, [vp, vp]),
            "dmg_closure_push_dev": (i32, [vp, i32, u32, u32, vp, vp, pu64]),
            "dmg_closure_pull_dev": (i32, [vp, i32, u32, u32, vp, vp, vp, pu64]),
            "dmg_closure_push_async": (i32, [vp, i32, u32, u32, vp, vp, vp, vp]),
            "dmg_closure_pull_async": (i32, [vp, i32, u32, u32, vp, vp, vp, vp, vp]),
            "dmg_closure_merge_async": (i32, [vp, u32, u32, u32, vp, u32, u32, vp, vp, vp, vp, vp]),
            "dmg_closure_merge_dev": (i32, [vp, u32, u32, u32, vp, u32, u32, vp, vp, vp, pu64]),
            "dmg_entry_points": (i32, [vp, vp, u32, C.POINTER(u32)]),
            "dmg_is_entry_point_batch": (i32, [vp, u64, vp, vp]),
            "dmg_search_batch": (i32, [vp, i32, u32, u32, u64, vp, vp, vp, vp, u64, pu64]),
            "dmg_lookup_batch": (i32, [vp, i32, u32, u32, u64, vp, vp, vp]),
            "dmg_query_batch": (i32, [vp, u64, vp, vp, vp, vp, vp, u64, pu64]),
            "dmg_graph_set_methods": (i32, [vp, vp, vp, vp]),
            "dmg_query_batch_meta": (i32, [vp, u64, vp, vp, vp, vp, vp, vp, u64, pu64]),
            "dmg_graph_set_class_types": (i32, [vp, vp]),
            "dmg_graph_set_method_flags": (i32, [vp, vp]),
            "dmg_last_call_stats": (i32, [vp, C.POINTER(CallStats)]),
            "dmg_profile_enable": (i32, [vp, i32]),
            "dmg_profile_reset": (i32, [vp]),
            "dmg_profile_count": (i32, [vp]),
            "dmg_profile_get": (i32, [vp, i32, C.POINTER(C.c_char_p), pu64, C.POINTER(C.c_double)]),
            "dmg_device_alloc": (i32, [vp, u64, C.POINTER(vp)]),
            "dmg_device_free": (i32, [vp, vp]),
            "dmg_copy_to_device": (i32, [vp, vp, vp, u64]),
            "dmg_copy_to_host": (i32, [vp, vp, vp, u64]),
            "dmg_pinned_alloc": (i32, [vp, u64, C.POINTER(vp)]),
            "dmg_pinned_free": (i32, [vp, vp]),
            "dmg_flush_l2": (i32, [vp]),
        }
        for name, (res, args) in sig.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


def bits_decode(planes, plane_words, win, counts, bits, out=None, threads=0):
    """dmg_bits_decode: bit rows -> (off u64[nq+1], ids u32[], hop u8[]) on the host."""
    nq = counts.size
    win = np.ascontiguousarray(win, np.uint32)
    counts = np.ascontiguousarray(counts, np.uint32)
    bits = np.ascontiguousarray(bits, np.uint32)
    need = C.c_uint64(0)
    if out is None:
        total = int(counts.sum(dtype=np.uint64))
        off = np.zeros(nq + 1, np.uint64)
        ids, hop = np.zeros(max(total, 1), np.uint32), np.zeros(max(total, 1), np.uint8)
    else:
        off, ids, hop = out
    _check(lib().dmg_bits_decode(planes, plane_words, nq, win.ctypes.data, counts.ctypes.data, bits.ctypes.data,
                                 off.ctypes.data, ids.ctypes.data, hop.ctypes.data, ids.size, C.byref(need), threads))
    return off, ids[:need.value], hop[:need.value]


def device_count():
    return lib().dmg_device_count()


def _check(rc, allow=()):
    if rc != OK and rc not in allow:
        raise DmgError(rc, lib().dmg_last_error().decode("utf-8", "replace"))
    return rc


def _arr(a, dt):
    a = np.ascontiguousarray(a, dtype=dt)
    return a


def _ptr(a):
    return a.ctypes.data if a is not None and a.size else None


class Engine:
    def __init__(self, device=0, scratch_bytes=0):
        self.L = lib()
        self.h = C.c_void_p()
        opts = EngineOpts(device, 0, scratch_bytes)
        _check(self.L.dmg_engine_create(C.byref(opts), C.byref(self.h)))
        self.device = device

    def close(self):
        if self.h:
            self.L.dmg_engine_destroy(self.h)
            self.h = C.c_void_p()

    def last_stats(self):
        s = CallStats()
        _check(self.L.dmg_last_call_stats(self.h, C.byref(s)))
        return {f: getattr(s, f) for f, _ in CallStats._fields_ if f != "reserved"}

    def profile(self, on=True):
        _check(self.L.dmg_profile_enable(self.h, int(on)))

    def profile_reset(self):
        _check(self.L.dmg_profile_reset(self.h))

    def profile_table(self):
        out = {}
        for i in range(self.L.dmg_profile_count(self.h)):
            name, n, ms = C.c_char_p(), C.c_uint64(), C.c_double()
            _check(self.L.dmg_profile_get(self.h, i, C.byref(name), C.byref(n), C.byref(ms)))
            if n.value:
                out[name.value.decode()] = (n.value, ms.value)
        return out

    def flush_l2(self):
        _check(self.L.dmg_flush_l2(self.h))

    # -- device / pinned buffers ---------------------------------------------------------------
    def device_alloc(self, nbytes):
        p = C.c_void_p()
        _check(self.L.dmg_device_alloc(self.h, int(nbytes), C.byref(p)))
        return p.value

    def device_free(self, p):
        _check(self.L.dmg_device_free(self.h, C.c_void_p(p)))

    def to_device(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.device_alloc(max(arr.nbytes, 16))
        if arr.nbytes:
            _check(self.L.dmg_copy_to_device(self.h, C.c_void_p(p), arr.ctypes.data, arr.nbytes))
        return p

    def to_host(self, p, dtype, count):
        out = np.empty(count, dtype=dtype)
        if out.nbytes:
            _check(self.L.dmg_copy_to_host(self.h, out.ctypes.data, C.c_void_p(p), out.nbytes))
        return out

    def pinned(self, dtype, count):
        """numpy array over page-locked host memory (kept alive by the engine until freed)."""
        p = C.c_void_p()
        nbytes = max(1, int(count)) * np.dtype(dtype).itemsize
        _check(self.L.dmg_pinned_alloc(self.h, nbytes, C.byref(p)))
        buf = (C.c_uint8 * nbytes).from_address(p.value)
        a = np.frombuffer(buf, dtype=dtype, count=int(count))
        return a, p.value

    def pinned_free(self, p):
        _check(self.L.dmg_pinned_free(self.h, C.c_void_p(p)))


class Graph:
    """One immutable device graph (dmg_graph handle)."""

    def __init__(self, engine, n_vertices, src, dst, entry=(), part_off=None, name_off=None, name_bytes=None,
                 lower_bytes=None):
        self.eng = engine
        self.L = engine.L
        src, dst = _arr(src, np.uint32), _arr(dst, np.uint32)
        entry = _arr(entry, np.uint32)
        keep = [src, dst, entry]
        d = GraphDesc()
        d.n_vertices = int(n_vertices)
        d.n_edges = int(src.size)
        d.edge_src, d.edge_dst = _ptr(src), _ptr(dst)
        d.n_entry = int(entry.size)
        d.entry_ids = _ptr(entry)
        if part_off is not None:
            po = _arr(part_off, np.uint32)
            keep.append(po)
            d.n_parts = int(po.size - 1)
            d.part_off = _ptr(po)
        if name_off is not None:
            no = _arr(name_off, np.uint64)
            nb = _arr(name_bytes, np.uint8)
            keep += [no, nb]
            d.name_off = no.ctypes.data
            d.name_bytes = nb.ctypes.data if nb.size else None
            if lower_bytes is not None:
                lb = _arr(lower_bytes, np.uint8)
                keep.append(lb)
                d.lower_bytes = lb.ctypes.data if lb.size else None
        self.h = C.c_void_p()
        _check(self.L.dmg_graph_create(engine.h, C.byref(d), C.byref(self.h)))
        info = GraphInfo()
        _check(self.L.dmg_graph_info_get(self.h, C.byref(info)))
        self.info = {f: getattr(info, f) for f, _ in GraphInfo._fields_}
        self.V = self.info["n_vertices"]

    def release(self):
        if self.h:
            self.L.dmg_graph_release(self.h)
            self.h = C.c_void_p()

    def save(self, path):
        _check(self.L.dmg_graph_save(self.h, path.encode()))

    @classmethod
    def load(cls, engine, path):
        h = C.c_void_p()
        _check(engine.L.dmg_graph_load(engine.h, path.encode(), C.byref(h)))
        g = cls.__new__(cls)
        g.eng, g.L, g.h = engine, engine.L, h
        info = GraphInfo()
        _check(g.L.dmg_graph_info_get(g.h, C.byref(info)))
        g.info = {f: getattr(info, f) for f, _ in GraphInfo._fields_}
        g.V = g.info["n_vertices"]
        return g

    # -- T0/T1 ---------------------------------------------------------------------------------
    def adjacency_batch(self, direction, ids, cap=None, out=None):
        ids = _arr(ids, np.uint32)
        n = ids.size
        need = C.c_uint64(0)
        if out is not None:                      # caller buffers (e.g. pinned): (off u64[n+1], ids u32[cap])
            off, o_ids = out
            _check(self.L.dmg_adjacency_batch(self.h, direction, n, _ptr(ids), off.ctypes.data, o_ids.ctypes.data,
                                              o_ids.size, C.byref(need)))
            return off, o_ids[:need.value]
        off = np.zeros(n + 1, np.uint64)
        if cap is None:
            rc = self.L.dmg_adjacency_batch(self.h, direction, n, _ptr(ids), off.ctypes.data, None, 0, C.byref(need))
            _check(rc, allow=(ECAP,))
            cap = need.value
        out = np.zeros(max(1, cap), np.uint32)
        _check(self.L.dmg_adjacency_batch(self.h, direction, n, _ptr(ids), off.ctypes.data, out.ctypes.data, cap,
                                          C.byref(need)))
        return off, out[:need.value]

    # -- T2/T3/T4 ------------------------------------------------------------------------------
    def neighbors_batch(self, mode, k, exclude_sources, q_off, q_src, cap=None, out=None):
        q_off, q_src = _arr(q_off, np.uint32), _arr(q_src, np.uint32)
        nq = q_off.size - 1
        need = C.c_uint64(0)
        if out is not None:
            off, ids, hop = out
            cap = ids.size
        else:
            off = np.zeros(nq + 1, np.uint64)
            if cap is None:
                rc = self.L.dmg_neighbors_batch(self.h, mode, k, int(exclude_sources), nq, _ptr(q_off), _ptr(q_src),
                                                off.ctypes.data, None, None, 0, C.byref(need))
                _check(rc, allow=(ECAP,))
                cap = need.value
            ids = np.zeros(max(1, cap), np.uint32)
            hop = np.zeros(max(1, cap), np.uint8)
        _check(self.L.dmg_neighbors_batch(self.h, mode, k, int(exclude_sources), nq, _ptr(q_off), _ptr(q_src),
                                          off.ctypes.data, ids.ctypes.data, hop.ctypes.data, cap, C.byref(need)))
        return off, ids[:need.value], hop[:need.value]

    def neighbors_batch_dev(self, mode, k, exclude_sources, nq, d_q_off, d_q_src, d_out_off, d_out_ids, d_out_hop,
                            cap):
        need = C.c_uint64(0)
        rc = self.L.dmg_neighbors_batch_dev(self.h, mode, k, int(exclude_sources), nq, C.c_void_p(d_q_off),
                                            C.c_void_p(d_q_src), C.c_void_p(d_out_off), C.c_void_p(d_out_ids),
                                            C.c_void_p(d_out_hop), cap, C.byref(need))
        _check(rc, allow=(ECAP,))
        return rc, need.value

    def neighbors_batch_rows(self, mode, k, exclude_sources, q_off, q_src, out=None):
        """Dense rows: returns (stride, win[nq,2], rows[nq,stride]) -- one hop byte per project vertex."""
        q_off, q_src = _arr(q_off, np.uint32), _arr(q_src, np.uint32)
        nq = q_off.size - 1
        stride, need = C.c_uint32(0), C.c_uint64(0)
        rc = self.L.dmg_neighbors_batch_rows(self.h, mode, k, int(exclude_sources), nq, _ptr(q_off), _ptr(q_src),
                                             C.byref(stride), None, None, 0, C.byref(need))
        _check(rc, allow=(ECAP,))
        if out is None:
            win = np.zeros((max(nq, 1), 2), np.uint32)
            rows = np.zeros(max(need.value, 1), np.uint8)
        else:
            win, rows = out
        _check(self.L.dmg_neighbors_batch_rows(self.h, mode, k, int(exclude_sources), nq, _ptr(q_off), _ptr(q_src),
                                               C.byref(stride), win.ctypes.data, rows.ctypes.data, rows.size,
                                               C.byref(need)))
        return stride.value, win[:nq], rows[:need.value].reshape(nq, stride.value) if nq else rows[:0]

    def neighbors_batch_bits(self, mode, k, exclude_sources, q_off, q_src, out=None):
        """Bit rows: returns (planes, plane_words, win[nq,2], counts[nq], bits[nq, planes, plane_words])."""
        q_off, q_src = _arr(q_off, np.uint32), _arr(q_src, np.uint32)
        nq = q_off.size - 1
        planes, words, need = C.c_uint32(0), C.c_uint32(0), C.c_uint64(0)
        rc = self.L.dmg_neighbors_batch_bits(self.h, mode, k, int(exclude_sources), nq, _ptr(q_off), _ptr(q_src),
                                             C.byref(planes), C.byref(words), None, None, None, 0, C.byref(need))
        _check(rc, allow=(ECAP,))
        if out is None:
            win = np.zeros((max(nq, 1), 2), np.uint32)
            counts = np.zeros(max(nq, 1), np.uint32)
            bits = np.zeros(max(need.value // 4, 1), np.uint32)
        else:
            win, counts, bits = out
        _check(self.L.dmg_neighbors_batch_bits(self.h, mode, k, int(exclude_sources), nq, _ptr(q_off), _ptr(q_src),
                                               C.byref(planes), C.byref(words), win.ctypes.data, counts.ctypes.data,
                                               bits.ctypes.data, bits.size * 4, C.byref(need)))
        n_words = need.value // 4
        return (planes.value, words.value, win[:nq], counts[:nq],
                bits[:n_words].reshape(nq, planes.value, words.value) if nq else bits[:0])

    def neighbors_batch_dev_static(self, mode, k, exclude_sources, nq, d_q_off, d_q_src, d_out_off, d_out_ids,
                                   d_out_hop, cap):
        """dmg_neighbors_batch_dev_static: same results, scratch in the library's static array."""
        need = C.c_uint64(0)
        rc = self.L.dmg_neighbors_batch_dev_static(self.h, mode, k, int(exclude_sources), nq, C.c_void_p(d_q_off),
                                                   C.c_void_p(d_q_src), C.c_void_p(d_out_off),
                                                   C.c_void_p(d_out_ids), C.c_void_p(d_out_hop), cap,
                                                   C.byref(need))
        _check(rc, allow=(ECAP,))
        return rc, need.value

    # -- T5 ------------------------------------------------------------------------------------
    def part(self, rank):
        """The per-device handle of `rank` as a Graph (borrowed: do not release it), e.g. to set the method
        catalog on every replica."""
        h = C.c_void_p()
        _check(self.L.dmg_dist_graph_part(self.h, rank, C.byref(h), None, None))
        g = Graph.__new__(Graph)
        g.eng, g.L, g.h = None, self.L, h
        info = GraphInfo()
        _check(self.L.dmg_graph_info_get(h, C.byref(info)))
        g.info = {f: getattr(info, f) for f, _ in GraphInfo._fields_}
        g.V = g.info["n_vertices"]
        return g

    def query_batch_packed(self, ops, text, out=None):
        """dmg_dist_query_batch: as Graph.query_batch_packed, the ops sharded over the replicas."""
        ops = np.ascontiguousarray(ops, dtype=Graph.QUERY_OP_DTYPE)
        tb = np.ascontiguousarray(text, dtype=np.uint8)
        if tb.size == 0:
            tb = np.zeros(1, np.uint8)
        need = C.c_uint64(0)
        if out is None:
            off = np.zeros(ops.size + 1, np.uint64)
            rc = self.L.dmg_dist_query_batch(self.h, ops.size, ops.ctypes.data, tb.ctypes.data, off.ctypes.data, None,
                                             None, 0, C.byref(need))
            _check(rc, allow=(ECAP,))
            return need.value
        off, ids, aux = out
        _check(self.L.dmg_dist_query_batch(self.h, ops.size, ops.ctypes.data, tb.ctypes.data, off.ctypes.data,
                                           ids.ctypes.data, aux.ctypes.data, ids.size, C.byref(need)))
        return need.value

    def analysis_order(self):
        """Exact analysisOrder of a partitioned graph (dmg_dist_analysis_order)."""
        order = np.zeros(max(1, self.V), np.uint32)
        _check(self.L.dmg_dist_analysis_order(self.h, order.ctypes.data))
        return order[:self.V]

    def closure(self, direction, sources):
        s = _arr(sources, np.uint32)
        level = np.zeros(max(1, self.V), np.uint32)
        reached = C.c_uint64(0)
        _check(self.L.dmg_closure(self.h, direction, s.size, _ptr(s), level.ctypes.data, C.byref(reached)))
        return level[:self.V], reached.value

    def closure_push_dev(self, direction, v_lo, v_hi, d_frontier, d_contrib):
        e = C.c_uint64(0)
        _check(self.L.dmg_closure_push_dev(self.h, direction, v_lo, v_hi, C.c_void_p(d_frontier),
                                           C.c_void_p(d_contrib), C.byref(e)))
        return e.value

    def closure_pull_dev(self, direction, v_lo, v_hi, d_frontier, d_visited, d_next_slice):
        e = C.c_uint64(0)
        _check(self.L.dmg_closure_pull_dev(self.h, direction, v_lo, v_hi, C.c_void_p(d_frontier),
                                           C.c_void_p(d_visited), C.c_void_p(d_next_slice), C.byref(e)))
        return e.value

    def closure_merge_dev(self, v_lo, v_hi, lvl, d_gathered, n_gathered, stride_words, d_visited, d_next, d_level):
        n = C.c_uint64(0)
        _check(self.L.dmg_closure_merge_dev(self.h, v_lo, v_hi, lvl, C.c_void_p(d_gathered), n_gathered,
                                            stride_words, C.c_void_p(d_visited), C.c_void_p(d_next),
                                            C.c_void_p(d_level), C.byref(n)))
        return n.value

    def analysis_order(self):
        order = np.zeros(max(1, self.V), np.uint32)
        _check(self.L.dmg_analysis_order(self.h, order.ctypes.data))
        return order[:self.V]

    def entry_points(self):
        need = C.c_uint32(0)
        out = np.zeros(max(1, self.info["n_entry"]), np.uint32)
        _check(self.L.dmg_entry_points(self.h, out.ctypes.data, out.size, C.byref(need)))
        return out[:need.value]

    def is_entry_point_batch(self, ids):
        ids = _arr(ids, np.uint32)
        out = np.zeros(max(1, ids.size), np.uint8)
        _check(self.L.dmg_is_entry_point_batch(self.h, ids.size, _ptr(ids), out.ctypes.data))
        return out[:ids.size]

    def lookup_batch(self, kind, texts, v_lo=0, v_hi=None):
        """texts: list of bytes. Returns uint32 ids, 0xFFFFFFFF = not found."""
        n = len(texts)
        t_off = np.zeros(n + 1, np.uint32)
        if n:
            np.cumsum([len(t) for t in texts], out=t_off[1:])
        tb = np.frombuffer(b"".join(texts) + b"\0", dtype=np.uint8).copy()
        out = np.zeros(max(n, 1), np.uint32)
        _check(self.L.dmg_lookup_batch(self.h, kind, v_lo, self.V if v_hi is None else v_hi, n, t_off.ctypes.data,
                                       tb.ctypes.data, out.ctypes.data))
        return out[:n]

    # -- Q1 ------------------------------------------------------------------------------------
    def set_methods(self, methods_per_vertex):
        """methods_per_vertex: list (len V) of lists of method names; lower-cased here (ASCII)."""
        mo = np.zeros(self.V + 1, np.uint32)
        names = []
        for v, ms in enumerate(methods_per_vertex):
            names.extend(ms)
            mo[v + 1] = len(names)
        enc = [m.lower().encode("utf-8") for m in names]
        no = np.zeros(len(enc) + 1, np.uint32)
        if enc:
            np.cumsum([len(b) for b in enc], out=no[1:])
        lb = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
        _check(self.L.dmg_graph_set_methods(self.h, mo.ctypes.data, no.ctypes.data, lb.ctypes.data))

    def set_method_names(self, methods_per_vertex):
        """Method names as written (for dmg_stacktrace_batch); same catalog order as set_methods."""
        enc = [m.encode("utf-8") for ms in methods_per_vertex for m in ms]
        no = np.zeros(len(enc) + 1, np.uint32)
        if enc:
            np.cumsum([len(b) for b in enc], out=no[1:])
        b = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
        _check(self.L.dmg_graph_set_method_names(self.h, no.ctypes.data, b.ctypes.data))

    def stacktrace_batch(self, traces, k=1):
        """traces: list of lists of (class_name, method_name or None).  Returns (status u8[frames],
        vertex u32[frames], method u32[frames], off u64[n+1], ids, hop)."""
        frames = [f for t in traces for f in t]
        n, nf = len(traces), len(frames)
        t_off = np.zeros(n + 1, np.uint32)
        if n:
            np.cumsum([len(t) for t in traces], out=t_off[1:])

        def pack(texts):
            off = np.zeros(len(texts) + 1, np.uint32)
            if texts:
                np.cumsum([len(x) for x in texts], out=off[1:])
            return off, np.frombuffer(b"".join(texts) + b"\0", dtype=np.uint8).copy()

        cn_off, cn = pack([c.encode("utf-8") for c, _m in frames])
        with_m = any(m is not None for _c, m in frames)
        mn_off, mn = pack([(m or "").encode("utf-8") for _c, m in frames])
        status = np.zeros(max(nf, 1), np.uint8)
        vertex = np.zeros(max(nf, 1), np.uint32)
        method = np.zeros(max(nf, 1), np.uint32)
        off = np.zeros(n + 1, np.uint64)
        need = C.c_uint64(0)
        args = (self.h, k, n, t_off.ctypes.data, cn_off.ctypes.data, cn.ctypes.data,
                mn_off.ctypes.data if with_m else None, mn.ctypes.data if with_m else None, status.ctypes.data,
                vertex.ctypes.data, method.ctypes.data, off.ctypes.data)
        rc = self.L.dmg_stacktrace_batch(*args, None, None, 0, C.byref(need))
        _check(rc, allow=(ECAP,))
        cap = need.value
        ids, hop = np.zeros(max(cap, 1), np.uint32), np.zeros(max(cap, 1), np.uint8)
        _check(self.L.dmg_stacktrace_batch(*args, ids.ctypes.data, hop.ctypes.data, cap, C.byref(need)))
        return status[:nf], vertex[:nf], method[:nf], off, ids[:need.value], hop[:need.value]

    def set_methods_packed(self, method_off, name_off, lower_bytes):
        """Array form of set_methods for large catalogs (u32 offsets, lower-cased bytes)."""
        mo = np.ascontiguousarray(method_off, dtype=np.uint32)
        no = np.ascontiguousarray(name_off, dtype=np.uint32)
        lb = np.ascontiguousarray(lower_bytes, dtype=np.uint8)
        assert mo.size == self.V + 1 and no.size == int(mo[-1]) + 1
        _check(self.L.dmg_graph_set_methods(self.h, mo.ctypes.data, no.ctypes.data, lb.ctypes.data))

    def set_method_flags(self, flags):
        f = np.ascontiguousarray(flags, dtype=np.uint8)
        _check(self.L.dmg_graph_set_method_flags(self.h, f.ctypes.data if f.size else None))

    def set_class_types(self, codes):
        c = np.ascontiguousarray(codes, dtype=np.uint8)
        assert c.size == self.V
        _check(self.L.dmg_graph_set_class_types(self.h, c.ctypes.data if c.size else None))

    def query_batch_meta(self, ops):
        """query_batch with the decoration column: rows of (ids, aux, meta)."""
        rows = self.query_batch(ops, with_meta=True)
        return rows

    def query_batch(self, ops, with_meta=False):
        """ops: list of dicts {op, vertex?, mode?, k?, exclude?, text?(bytes), project?}; returns rows of (id, aux)."""
        n = len(ops)
        arr = (QueryOp * max(n, 1))()
        text = bytearray()
        for i, o in enumerate(ops):
            a = arr[i]
            a.opcode = o["op"]
            a.vertex = o.get("vertex", 0xFFFFFFFF)
            a.mode, a.k, a.exclude_sources = o.get("mode", 0), o.get("k", 1), int(o.get("exclude", 0))
            a.project = o.get("project", 0)
            t = o.get("text", b"")
            a.text_begin = len(text)
            text += t
            a.text_end = len(text)
        tb = np.frombuffer(bytes(text) + b"\0", dtype=np.uint8).copy()
        off = np.zeros(n + 1, np.uint64)
        need = C.c_uint64(0)
        rc = self.L.dmg_query_batch(self.h, n, arr, tb.ctypes.data, off.ctypes.data, None, None, 0, C.byref(need))
        _check(rc, allow=(ECAP,))
        cap = need.value
        ids, aux = np.zeros(max(1, cap), np.uint32), np.zeros(max(1, cap), np.uint8)
        if with_meta:
            meta = np.zeros(max(1, cap), np.uint8)
            _check(self.L.dmg_query_batch_meta(self.h, n, arr, tb.ctypes.data, off.ctypes.data, ids.ctypes.data,
                                               aux.ctypes.data, meta.ctypes.data, cap, C.byref(need)))
            return [(ids[int(off[i]):int(off[i + 1])].tolist(), aux[int(off[i]):int(off[i + 1])].tolist(),
                     meta[int(off[i]):int(off[i + 1])].tolist()) for i in range(n)]
        _check(self.L.dmg_query_batch(self.h, n, arr, tb.ctypes.data, off.ctypes.data, ids.ctypes.data,
                                      aux.ctypes.data, cap, C.byref(need)))
        return [(ids[int(off[i]):int(off[i + 1])].tolist(), aux[int(off[i]):int(off[i + 1])].tolist())
                for i in range(n)]

    QUERY_OP_DTYPE = np.dtype([("opcode", "<u4"), ("vertex", "<u4"), ("mode", "<u4"), ("k", "<u4"),
                               ("exclude_sources", "<u4"), ("text_begin", "<u4"), ("text_end", "<u4"),
                               ("project", "<u4")])

    def query_batch_packed(self, ops, text, out=None):
        """ops: structured array of QUERY_OP_DTYPE (== dmg_query_op), text: uint8 array.
        out=None sizes the result (returns needed); out=(off u64[n+1], ids u32[cap], aux u8[cap])
        runs the batch into those buffers and returns needed."""
        ops = np.ascontiguousarray(ops, dtype=self.QUERY_OP_DTYPE)
        tb = np.ascontiguousarray(text, dtype=np.uint8)
        if tb.size == 0:
            tb = np.zeros(1, np.uint8)
        n = ops.size
        need = C.c_uint64(0)
        if out is None:
            off = np.zeros(n + 1, np.uint64)
            rc = self.L.dmg_query_batch(self.h, n, ops.ctypes.data, tb.ctypes.data, off.ctypes.data, None, None, 0,
                                        C.byref(need))
            _check(rc, allow=(ECAP,))
            return need.value
        off, ids, aux = out
        _check(self.L.dmg_query_batch(self.h, n, ops.ctypes.data, tb.ctypes.data, off.ctypes.data, ids.ctypes.data,
                                      aux.ctypes.data, ids.size, C.byref(need)))
        return need.value

    # -- S1/S2 ---------------------------------------------------------------------------------
    def search_batch(self, kind, queries, v_lo=0, v_hi=None, cap=None, out=None):
        """queries: list of bytes, already lower-cased by the host, or a packed (q_off u32[nq+1], q_bytes u8[])
        pair.  out=(off u64[nq+1], ids u32[cap]) runs into caller buffers (e.g. pinned ones)."""
        if isinstance(queries, tuple):
            q_off, qb = queries
            nq = q_off.size - 1
        else:
            nq = len(queries)
            q_off = np.zeros(nq + 1, np.uint32)
            if nq:
                np.cumsum([len(q) for q in queries], out=q_off[1:])
            qb = np.frombuffer(b"".join(queries) + b"\0", dtype=np.uint8).copy()
        if v_hi is None:
            v_hi = self.V
        need = C.c_uint64(0)
        if out is not None:
            off, ids = out
            _check(self.L.dmg_search_batch(self.h, kind, v_lo, v_hi, nq, q_off.ctypes.data, qb.ctypes.data,
                                           off.ctypes.data, ids.ctypes.data, ids.size, C.byref(need)))
            return off, ids[:need.value]
        off = np.zeros(nq + 1, np.uint64)
        if cap is None:
            rc = self.L.dmg_search_batch(self.h, kind, v_lo, v_hi, nq, q_off.ctypes.data, qb.ctypes.data,
                                         off.ctypes.data, None, 0, C.byref(need))
            _check(rc, allow=(ECAP,))
            cap = need.value
        ids = np.zeros(max(1, cap), np.uint32)
        _check(self.L.dmg_search_batch(self.h, kind, v_lo, v_hi, nq, q_off.ctypes.data, qb.ctypes.data,
                                       off.ctypes.data, ids.ctypes.data, cap, C.byref(need)))
        return off, ids[:need.value]


class Dist:
    """dmg_dist: one engine per device of the box inside this process (include/dmgraph.h)."""

    def __init__(self, n, devices=None):
        self.L = lib()
        self.h = C.c_void_p()
        dv = (C.c_int32 * n)(*devices) if devices is not None else None
        _check(self.L.dmg_dist_create(dv, n, C.byref(self.h)))
        self.n = n

    def close(self):
        if self.h:
            self.L.dmg_dist_destroy(self.h)
            self.h = C.c_void_p()


class DistGraph:
    """A graph on every device of a Dist: replicated (query batches) or vertex-range partitioned (closure)."""

    def __init__(self, dist, n_vertices, src, dst, entry=(), part_off=None, partitioned=False, name_off=None,
                 name_bytes=None, lower_bytes=None):
        self.dist, self.L, self.V = dist, dist.L, int(n_vertices)
        src, dst, entry = _arr(src, np.uint32), _arr(dst, np.uint32), _arr(entry, np.uint32)
        d = GraphDesc()
        d.n_vertices, d.n_edges, d.n_entry = self.V, int(src.size), int(entry.size)
        d.edge_src, d.edge_dst, d.entry_ids = _ptr(src), _ptr(dst), _ptr(entry)
        if part_off is not None:
            po = _arr(part_off, np.uint32)
            d.n_parts, d.part_off = int(po.size - 1), _ptr(po)
        if name_off is not None:
            no, nb = _arr(name_off, np.uint64), _arr(name_bytes, np.uint8)
            d.name_off, d.name_bytes = no.ctypes.data, (nb.ctypes.data if nb.size else None)
            if lower_bytes is not None:
                lb = _arr(lower_bytes, np.uint8)
                d.lower_bytes = lb.ctypes.data if lb.size else None
        self.h = C.c_void_p()
        _check(self.L.dmg_dist_graph_create(dist.h, C.byref(d), int(partitioned), C.byref(self.h)))

    def release(self):
        if self.h:
            self.L.dmg_dist_graph_release(self.h)
            self.h = C.c_void_p()

    def neighbors_batch(self, mode, k, exclude_sources, q_off, q_src, out=None):
        q_off, q_src = _arr(q_off, np.uint32), _arr(q_src, np.uint32)
        nq = q_off.size - 1
        need = C.c_uint64(0)
        if out is None:
            off = np.zeros(nq + 1, np.uint64)
            rc = self.L.dmg_dist_neighbors_batch(self.h, mode, k, int(exclude_sources), nq, _ptr(q_off), _ptr(q_src),
                                                 off.ctypes.data, None, None, 0, C.byref(need))
            _check(rc, allow=(ECAP,))
            ids, hop = np.zeros(max(1, need.value), np.uint32), np.zeros(max(1, need.value), np.uint8)
        else:
            off, ids, hop = out
        _check(self.L.dmg_dist_neighbors_batch(self.h, mode, k, int(exclude_sources), nq, _ptr(q_off), _ptr(q_src),
                                               off.ctypes.data, ids.ctypes.data, hop.ctypes.data, ids.size,
                                               C.byref(need)))
        return off, ids[:need.value], hop[:need.value]

    def part(self, rank):
        """The per-device handle of `rank` as a Graph (borrowed: do not release it), e.g. to set the method
        catalog on every replica."""
        h = C.c_void_p()
        _check(self.L.dmg_dist_graph_part(self.h, rank, C.byref(h), None, None))
        g = Graph.__new__(Graph)
        g.eng, g.L, g.h = None, self.L, h
        info = GraphInfo()
        _check(self.L.dmg_graph_info_get(h, C.byref(info)))
        g.info = {f: getattr(info, f) for f, _ in GraphInfo._fields_}
        g.V = g.info["n_vertices"]
        return g

    def query_batch_packed(self, ops, text, out=None):
        """dmg_dist_query_batch: as Graph.query_batch_packed, the ops sharded over the replicas."""
        ops = np.ascontiguousarray(ops, dtype=Graph.QUERY_OP_DTYPE)
        tb = np.ascontiguousarray(text, dtype=np.uint8)
        if tb.size == 0:
            tb = np.zeros(1, np.uint8)
        need = C.c_uint64(0)
        if out is None:
            off = np.zeros(ops.size + 1, np.uint64)
            rc = self.L.dmg_dist_query_batch(self.h, ops.size, ops.ctypes.data, tb.ctypes.data, off.ctypes.data, None,
                                             None, 0, C.byref(need))
            _check(rc, allow=(ECAP,))
            return need.value
        off, ids, aux = out
        _check(self.L.dmg_dist_query_batch(self.h, ops.size, ops.ctypes.data, tb.ctypes.data, off.ctypes.data,
                                           ids.ctypes.data, aux.ctypes.data, ids.size, C.byref(need)))
        return need.value

    def analysis_order(self):
        """Exact analysisOrder of a partitioned graph (dmg_dist_analysis_order)."""
        order = np.zeros(max(1, self.V), np.uint32)
        _check(self.L.dmg_dist_analysis_order(self.h, order.ctypes.data))
        return order[:self.V]

    def closure(self, direction, sources):
        s = _arr(sources, np.uint32)
        level = np.zeros(max(1, self.V), np.uint32)
        reached, st = C.c_uint64(0), DistClosureStats()
        _check(self.L.dmg_dist_closure(self.h, direction, s.size, _ptr(s), level.ctypes.data, C.byref(reached),
                                       C.byref(st)))
        return level[:self.V], reached.value, {f: getattr(st, f) for f, _ in DistClosureStats._fields_
                                               if not f.startswith("reserved")}
