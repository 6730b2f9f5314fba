"""TEST INFRASTRUCTURE -- full-size result checks shared by tests/ and by bench.py's post-run parity
leg (the oracle as the checker, never as the thing measured).  Every function compares rows that
came out of libdmgraph.so with the CPU restatements of oracle.cc and raises AssertionError on the
first difference; the return value is the number of rows compared."""
import numpy as np

from . import oracle as ora


def check_khop_rows(o, mode, k, exclude, q_off, q_src, off, ids, hop, rows=None, threads=None):
    """Rows `rows` (indices into the batch; None = all) of a k-hop result against CsrOracle."""
    threads = threads or ora.hardware_threads()
    q_off = np.asarray(q_off, np.uint32)
    q_src = np.asarray(q_src, np.uint32)
    if rows is None:
        woff, wids, whop, _ = o.khop_batch(mode, k, exclude, q_off, q_src, threads=threads)
        assert np.array_equal(np.asarray(off), woff), "k-hop row offsets differ from the oracle"
        assert np.array_equal(np.asarray(ids[:int(woff[-1])]), wids), "k-hop ids differ from the oracle"
        assert np.array_equal(np.asarray(hop[:int(woff[-1])]), whop), "k-hop hops differ from the oracle"
        return q_off.size - 1
    rows = np.asarray(rows, np.int64)
    cnt = (q_off[rows + 1] - q_off[rows]).astype(np.int64)
    s_off = np.zeros(rows.size + 1, np.uint32)
    np.cumsum(cnt, out=s_off[1:])
    s_src = np.concatenate([q_src[int(q_off[r]):int(q_off[r + 1])] for r in rows]) if rows.size else q_src[:0]
    woff, wids, whop, _ = o.khop_batch(mode, k, exclude, s_off, s_src, threads=threads)
    for j, r in enumerate(rows):
        a, b = int(off[r]), int(off[r + 1])
        wa, wb = int(woff[j]), int(woff[j + 1])
        assert b - a == wb - wa, "k-hop row %d: %d results, oracle %d" % (r, b - a, wb - wa)
        assert np.array_equal(ids[a:b], wids[wa:wb]) and np.array_equal(hop[a:b], whop[wa:wb]), "k-hop row %d differs" % r
    return rows.size


def check_query_mix(o, ops, text, off, ids, aux, name_off, lower, mbytes, endpoint_methods, rows=None,
                    threads=None):
    """Rows of a dmg_query_batch result (BASELINE config 4 mix) against the restatements: EVERY opcode
    is checked -- adjacency rows, overview, directed k-hop, entry points, endpoints, simple-name /
    containment resolution (GQS:431-453) and the method check (GQS:362-394) over the 10-per-vertex
    catalog of synth.method_catalog."""
    threads = threads or ora.hardware_threads()
    n = len(ops)
    rows = np.arange(n) if rows is None else np.asarray(rows, np.int64)
    opc = ops["opcode"][rows]
    vert = ops["vertex"][rows]
    V = o.V

    def row(i):
        a, b = int(off[i]), int(off[i + 1])
        return ids[a:b], aux[a:b]

    # adjacency-shaped ops, batched through the oracle
    for code, dirs in ((0, (0,)), (1, (1,)), (2, (0, 1))):
        sel = rows[opc == code]
        if not sel.size:
            continue
        want = [o.adjacency_batch(d, ops["vertex"][sel]) for d in dirs]
        for j, i in enumerate(sel):
            gi, ga = row(i)
            parts = [w[1][int(w[0][j]):int(w[0][j + 1])] for w in want]
            wi = np.concatenate(parts)
            wa = np.concatenate([np.full(p.size, d if code == 2 else 0, np.uint8) for d, p in enumerate(parts)])
            assert np.array_equal(gi, wi) and np.array_equal(ga, wa), "op %d (opcode %d) differs" % (i, code)
    # k-hop groups
    sel = rows[opc == 3]
    for key in {(int(ops["mode"][i]), int(ops["k"][i]), int(ops["exclude_sources"][i])) for i in sel}:
        grp = np.array([i for i in sel if (int(ops["mode"][i]), int(ops["k"][i]), int(ops["exclude_sources"][i])) == key])
        woff, wids, whop, _ = o.khop_batch(key[0], key[1], key[2], np.arange(grp.size + 1, dtype=np.uint32),
                                           ops["vertex"][grp], threads=threads)
        for j, i in enumerate(grp):
            gi, ga = row(i)
            assert np.array_equal(gi, wids[int(woff[j]):int(woff[j + 1])]), "k-hop op %d ids differ" % i
            assert np.array_equal(ga, whop[int(woff[j]):int(woff[j + 1])]), "k-hop op %d hops differ" % i
    # entry points / endpoints (single-project handle: the whole lists)
    ep = o.entry_points()
    for i in rows[opc == 4]:
        gi, ga = row(i)
        assert np.array_equal(gi, ep) and not ga.any(), "entrypoints op %d differs" % i
    for i in rows[opc == 7]:
        gi, ga = row(i)
        assert np.array_equal(gi, endpoint_methods) and not ga.any(), "endpoints op %d differs" % i
    # name resolution: first simple-name equality, else first containment
    sel = rows[opc == 5]
    if sel.size:
        qs = [bytes(text[int(ops["text_begin"][i]):int(ops["text_end"][i])]) for i in sel]
        o1, i1 = ora.search_batch(name_off, lower, qs, 1, threads=threads)
        miss = [j for j in range(sel.size) if o1[j + 1] == o1[j]]
        o2, i2 = ora.search_batch(name_off, lower, [qs[j] for j in miss], 2, threads=threads)
        second = {j: (int(i2[int(o2[m])]) if o2[m + 1] > o2[m] else None) for m, j in enumerate(miss)}
        for j, i in enumerate(sel):
            gi, ga = row(i)
            if o1[j + 1] > o1[j]:
                want = ([int(i1[int(o1[j])])], [2])
            elif second[j] is not None:
                want = ([second[j]], [3])
            else:
                want = ([], [])
            assert (gi.tolist(), ga.tolist()) == want, "resolve_name op %d: %r != %r" % (i, (gi.tolist(), ga.tolist()), want)
    # method check: first method of the vertex whose lower-cased name equals the text
    per = 10
    for i in rows[opc == 6]:
        gi, ga = row(i)
        v = int(ops["vertex"][i])
        t = bytes(text[int(ops["text_begin"][i]):int(ops["text_end"][i])])
        want = []
        if v < V:
            for j in range(per):
                if bytes(mbytes[(v * per + j) * 8:(v * per + j + 1) * 8]) == t:
                    want = [v * per + j]
                    break
        assert gi.tolist() == want, "method op %d differs" % i
    del vert
    return rows.size
