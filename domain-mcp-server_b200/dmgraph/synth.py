"""Seeded synthetic workloads of the shapes BASELINE.json names.

Everything here is deterministic numpy (PCG64 with the seeds SURVEY.md fixes) and produces
exactly what the Java host would hand the engine: a node-name list in `nodes` order, an edge
list in addDependency order, entry points, project (partition) ranges, and query batches.

  eco_graph(...)        C2: P disjoint project graphs (block-diagonal), log-normal out-degree,
                        Zipf(s=1) targets without replacement over a per-project permutation
  big_graph(...)        C4 / C5: one project, same generator, larger caps
  stacktrace_queries()  C2: random walks along out-edges from entry points, <=12 matched frames
  catalog_names()       C3: method-qualified identifiers
  search_queries()      C3: 50/40/10 % selectivity mix, already lower-cased
"""
import numpy as np

SEED_C2_GRAPH, SEED_C2_QUERIES = 20261106, 20261107
SEED_C3_NAMES, SEED_C3_QUERIES = 20261108, 20261109
SEED_C4, SEED_C5 = 20261110, 20261111

LAYERS = ["controller", "service", "domain", "repository", "dto"]
SUFFIX = {"controller": "Controller", "service": "Service", "domain": "", "repository": "Repository",
          "dto": "Dto"}
STEMS = ["Order", "User", "Payment", "Invoice", "Cart", "Catalog", "Shipment", "Account", "Ledger",
         "Inventory", "Pricing", "Promo", "Refund", "Session", "Token", "Audit", "Report", "Search",
         "Notification", "Webhook", "Tenant", "Billing", "Contract", "Customer", "Vendor", "Product",
         "Review", "Wishlist", "Address", "Tax"]
VERBS = ["create", "find", "update", "delete", "process", "validate", "list", "get", "save", "load",
         "sync", "publish", "resolve", "compute", "render", "parse"]
NOUNS = ["Order", "User", "Batch", "Page", "Item", "Total", "Status", "Event", "Record", "Entry",
         "Summary", "Detail", "Config", "Context", "Result", "Request"]
TS_KINDS = ["component", "service", "store", "hook", "util"]


def _degrees(rng, n, total, cap, sigma=0.7):
    """Out-degrees d_v = clamp(round(LogNormal(mu, sigma)), 0, cap) with sum exactly `total`."""
    cap = min(cap, n - 1)
    z = rng.standard_normal(n)
    lo, hi = -10.0, 10.0
    for _ in range(60):
        mu = 0.5 * (lo + hi)
        d = np.clip(np.rint(np.exp(mu + sigma * z)), 0, cap)
        if d.sum() < total:
            lo = mu
        else:
            hi = mu
    d = np.clip(np.rint(np.exp(hi + sigma * z)), 0, cap).astype(np.int64)
    diff = int(d.sum() - total)
    order = rng.permutation(n)
    i = 0
    while diff != 0:                         # nudge single vertices until the sum is exact
        v = order[i % n]
        i += 1
        if diff > 0 and d[v] > 0:
            d[v] -= 1
            diff -= 1
        elif diff < 0 and d[v] < cap:
            d[v] += 1
            diff += 1
    return d


def _zipf_targets(rng, n, deg):
    """For every vertex v draw deg[v] DISTINCT targets != v, Zipf(s=1) over a random ranking.

    Successive sampling without replacement, done as vectorised rejection: draw with
    replacement from the Zipf CDF, drop repeats and self-references keeping first
    occurrences, redraw the shortfall.  Returns (src, dst) in per-source draw order, sources
    ascending -- the order the host would call addDependency in.
    """
    perm = rng.permutation(n)                       # rank r -> vertex perm[r]
    cdf = np.cumsum(1.0 / np.arange(1, n + 1))
    cdf /= cdf[-1]
    need = deg.copy()
    got_src = [np.empty(0, np.int64)]
    got_dst = [np.empty(0, np.int64)]
    got_seq = [np.empty(0, np.int64)]
    seq_base = 0
    keys_seen = np.empty(0, np.int64)
    while need.sum() > 0:
        vs = np.repeat(np.arange(n), need * 2 + 2 * (need > 0))     # oversample 2x
        t = perm[np.searchsorted(cdf, rng.random(vs.size), side="left").clip(0, n - 1)]
        seq = seq_base + np.arange(vs.size)
        seq_base += vs.size
        ok = t != vs
        vs, t, seq = vs[ok], t[ok], seq[ok]
        key = vs * n + t
        # drop keys already accepted in earlier rounds
        if keys_seen.size:
            ok = ~np.isin(key, keys_seen)
            vs, t, seq, key = vs[ok], t[ok], seq[ok], key[ok]
        # first occurrence of each key within this round
        _u, first = np.unique(key, return_index=True)
        first.sort()
        vs, t, seq, key = vs[first], t[first], seq[first], key[first]
        # keep only the first `need[v]` per vertex (vs is non-decreasing)
        start = np.searchsorted(vs, np.arange(n), side="left")
        rank = np.arange(vs.size) - start[vs]
        ok = rank < need[vs]
        vs, t, seq, key = vs[ok], t[ok], seq[ok], key[ok]
        got_src.append(vs)
        got_dst.append(t)
        got_seq.append(seq)
        keys_seen = np.concatenate([keys_seen, key])
        need = need - np.bincount(vs, minlength=n)
    src = np.concatenate(got_src)
    dst = np.concatenate(got_dst)
    seq = np.concatenate(got_seq)
    o = np.lexsort((seq, src))
    return src[o], dst[o]


def project_names(p, n, rng):
    """60 % Java FQCNs, 20 % Go package paths, 20 % TS module ids; ASCII, mean ~ 45-50 B."""
    kind = rng.random(n)
    layer = rng.integers(0, len(LAYERS), n)
    stem = rng.integers(0, len(STEMS), n)
    tsk = rng.integers(0, len(TS_KINDS), n)
    out = []
    for i in range(n):
        L, S = LAYERS[layer[i]], STEMS[stem[i]]
        if kind[i] < 0.6:
            out.append("co.fanki.svc%d.%s.%s%d%s" % (p, L, S, i, SUFFIX[L]))
        elif kind[i] < 0.8:
            out.append("github.com/fanki/svc%d/internal/%s/%s%d" % (p, L, S.lower(), i))
        else:
            out.append("modules.svc%d.%s.%s%d.%s" % (p, S.lower(), S.lower(), i, TS_KINDS[tsk[i]]))
    return out


class SynthGraph:
    """What the host hands to dmg_graph_create: names, edges (insertion order), entries, parts."""

    def __init__(self, names, src, dst, entry, part_off):
        self.names = names                      # list[str] or None (graph-only workloads)
        self.src = np.ascontiguousarray(src, dtype=np.uint32)
        self.dst = np.ascontiguousarray(dst, dtype=np.uint32)
        self.entry = np.ascontiguousarray(entry, dtype=np.uint32)
        self.part_off = np.ascontiguousarray(part_off, dtype=np.uint32)
        self.n_vertices = int(part_off[-1])

    @property
    def n_edges(self):
        return int(self.src.size)


def eco_graph(n_projects=50, v_per=2000, e_per=40000, cap=200, entry_frac=0.02,
              seed=SEED_C2_GRAPH, with_names=True):
    """BASELINE config 2: `n_projects` disjoint project graphs concatenated (block diagonal)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    srcs, dsts, entries, names = [], [], [], []
    for p in range(n_projects):
        base = p * v_per
        deg = _degrees(rng, v_per, e_per, cap)
        s, d = _zipf_targets(rng, v_per, deg)
        srcs.append(s + base)
        dsts.append(d + base)
        ne = max(1, int(round(entry_frac * v_per)))
        entries.append(np.sort(rng.choice(v_per, ne, replace=False)) + base)
        if with_names:
            names.extend(project_names(p, v_per, rng))
    part_off = np.arange(n_projects + 1, dtype=np.int64) * v_per
    return SynthGraph(names if with_names else None, np.concatenate(srcs), np.concatenate(dsts),
                      np.concatenate(entries), part_off)


def big_graph(n_vertices, n_edges, cap, entry_frac, seed, with_names=False):
    """BASELINE configs 4 / 5: one project graph."""
    rng = np.random.Generator(np.random.PCG64(seed))
    deg = _degrees(rng, n_vertices, n_edges, cap)
    s, d = _zipf_targets(rng, n_vertices, deg)
    ne = max(1, int(round(entry_frac * n_vertices)))
    entry = np.sort(rng.choice(n_vertices, ne, replace=False))
    names = project_names(0, n_vertices, rng) if with_names else None
    return SynthGraph(names, s, d, entry, np.array([0, n_vertices]))


def big_graph_fast(n_vertices, n_edges, cap, entry_frac, seed, threads=None):
    """Config-5-sized graphs (20 M / 500 M): same out-degree law and Zipf(s=1) target ranking as
    big_graph, but targets are drawn WITH replacement in one vectorised pass (the exact
    without-replacement rejection loop needs several sorts of 10^9 keys).  Repeated (src, dst) pairs
    are legal input -- addDependency collapses them (ProjectGraph.java:182) and so does
    dmg_graph_create -- so the graph the engine holds has slightly fewer edges than drawn; the
    kept count is in dmg_graph_info.  Targets are drawn in fixed chunks of 8 M edges, chunk i from
    the i-th jump of the seed's PCG64 stream, on a thread pool (numpy releases the GIL): the
    result depends on the seed only, not on the thread count."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    rng = np.random.Generator(np.random.PCG64(seed))
    deg = _degrees(rng, n_vertices, n_edges, cap)
    perm = rng.permutation(n_vertices).astype(np.uint32)
    cdf = np.cumsum(1.0 / np.arange(1, n_vertices + 1))
    cdf /= cdf[-1]
    src = np.repeat(np.arange(n_vertices, dtype=np.uint32), deg)
    dst = np.empty(src.size, np.uint32)
    step = 8_000_000
    base = np.random.PCG64(seed ^ 0xD57)

    def fill(i):
        a, b = i * step, min(src.size, (i + 1) * step)
        r = np.random.Generator(base.jumped(i + 1))
        dst[a:b] = perm[np.searchsorted(cdf, r.random(b - a), side="left").clip(0, n_vertices - 1)]

    n_chunks = (src.size + step - 1) // step
    with ThreadPoolExecutor(max_workers=threads or min(16, os.cpu_count() or 1)) as ex:
        list(ex.map(fill, range(n_chunks)))
    ne = max(1, int(round(entry_frac * n_vertices)))
    entry = np.sort(rng.choice(n_vertices, ne, replace=False))
    return SynthGraph(None, src, dst, entry, np.array([0, n_vertices]))


def out_csr(g):
    """Forward CSR of a SynthGraph (host-side helper for the query generators)."""
    n = g.n_vertices
    cnt = np.bincount(g.src, minlength=n)
    ptr = np.zeros(n + 1, np.int64)
    np.cumsum(cnt, out=ptr[1:])
    o = np.argsort(g.src, kind="stable")
    return ptr, g.dst[o].astype(np.int64)


def stacktrace_queries(g, n_queries=65536, frames=12, seed=SEED_C2_QUERIES):
    """BASELINE config 2 query batch.

    Each trace picks a project uniformly, starts at a random entry point of it and walks
    `frames` steps along out-edges (restart at a random vertex of the project on a dead end).
    The 4 out-of-graph frames per trace (org.springframework...) are dropped by the host before
    the call and so never reach the engine.  Returns CSR-shaped (q_off, q_src); repeated
    classes within a trace are kept (the engine must cope, CodeContextService.java:1563-1568
    dedups them through a HashSet).
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    ptr, idx = out_csr(g)
    n_parts = g.part_off.size - 1
    part = rng.integers(0, n_parts, n_queries)
    lo = g.part_off[part].astype(np.int64)
    hi = g.part_off[part + 1].astype(np.int64)
    # entry points per partition
    e_part = np.searchsorted(g.part_off, g.entry, side="right") - 1
    e_order = np.argsort(e_part, kind="stable")
    e_sorted = g.entry[e_order].astype(np.int64)
    e_start = np.searchsorted(e_part[e_order], np.arange(n_parts), side="left")
    e_cnt = np.bincount(e_part, minlength=n_parts)
    pick = (rng.random(n_queries) * e_cnt[part]).astype(np.int64)
    cur = np.where(e_cnt[part] > 0, e_sorted[np.minimum(e_start[part] + pick, e_sorted.size - 1)],
                   lo + (rng.random(n_queries) * (hi - lo)).astype(np.int64))
    out = np.empty((n_queries, frames), np.int64)
    for f in range(frames):
        out[:, f] = cur
        deg = ptr[cur + 1] - ptr[cur]
        step = ptr[cur] + (rng.random(n_queries) * np.maximum(deg, 1)).astype(np.int64)
        nxt = idx[np.minimum(step, idx.size - 1)]
        restart = lo + (rng.random(n_queries) * (hi - lo)).astype(np.int64)
        cur = np.where(deg > 0, nxt, restart)
    q_off = (np.arange(n_queries + 1, dtype=np.int64) * frames).astype(np.uint32)
    return q_off, out.reshape(-1).astype(np.uint32)


def pack_names(names):
    """names -> (name_off u64[V+1], bytes u8[...]) UTF-8 arena."""
    enc = [s.encode("utf-8") for s in names]
    off = np.zeros(len(enc) + 1, np.uint64)
    np.cumsum([len(b) for b in enc], out=off[1:])
    return off, np.frombuffer(b"".join(enc), dtype=np.uint8).copy()


SYLLABLES = ["ba", "ce", "di", "fo", "gu", "ha", "je", "ki", "lo", "mu", "na", "pe", "qui", "ro", "su", "ta",
             "ve", "wi", "xo", "yu", "za", "bre", "cli", "dro", "fra", "gle", "pri", "sto", "tru", "vla",
             "sha", "che", "tho", "phi", "kra", "sno", "twi", "spe", "sca", "blu"]


def catalog_vocab(n_words=20000, seed=SEED_C3_NAMES):
    """Synthetic stem vocabulary: 3-syllable words, 6-9 letters, capitalised ("Brecliro")."""
    rng = np.random.Generator(np.random.PCG64(seed ^ 0x5EED))
    ns = len(SYLLABLES)
    seen, words = set(), []
    while len(words) < n_words:
        i, j, k = rng.integers(0, ns, 3)
        w = SYLLABLES[i] + SYLLABLES[j] + SYLLABLES[k]
        if 6 <= len(w) <= 10 and w not in seen:
            seen.add(w)
            words.append(w.capitalize())
    return words


def catalog_names(n, seed=SEED_C3_NAMES, n_projects=50, n_words=20000):
    """BASELINE config 3: method-qualified identifiers "{class}.{verb}{Noun}{n}" as a packed arena.

    The class stem comes from a 20,000-word vocabulary so that a vocabulary-token query has the
    ~1e-4 selectivity SURVEY.md 8d asks for.  Built directly as bytes: returns (name_off, bytes).
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    vocab = catalog_vocab(n_words, seed)
    chunks = []
    lens = np.empty(n, np.int64)
    step = 200000
    for base in range(0, n, step):
        m = min(step, n - base)
        p = rng.integers(0, n_projects, m)
        layer = rng.integers(0, len(LAYERS), m)
        stem = rng.integers(0, len(vocab), m)
        verb = rng.integers(0, len(VERBS), m)
        noun = rng.integers(0, len(NOUNS), m)
        cls = rng.integers(0, 2000, m)
        strs = ["co.fanki.svc%d.%s.%s%d%s.%s%s%d" % (p[i], LAYERS[layer[i]], vocab[stem[i]], cls[i],
                                                    SUFFIX[LAYERS[layer[i]]], VERBS[verb[i]],
                                                    NOUNS[noun[i]], base + i) for i in range(m)]
        enc = [s.encode("ascii") for s in strs]
        lens[base:base + m] = [len(b) for b in enc]
        chunks.append(b"".join(enc))
    off = np.zeros(n + 1, np.uint64)
    np.cumsum(lens, out=off[1:])
    return off, np.frombuffer(b"".join(chunks), dtype=np.uint8).copy()


def search_queries(name_off, name_bytes, n_queries=16384, seed=SEED_C3_QUERIES):
    """BASELINE config 3 queries, already lower-cased (the Java host lowers).

    50 % exact simple names of random catalog entries, 40 % 6-10 char vocabulary tokens,
    10 % 3-4 char substrings.  Returns a list of bytes objects.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    n = name_off.size - 1
    vocab = [w.lower() for w in catalog_vocab()]
    out = []
    kinds = rng.random(n_queries)
    for qi in range(n_queries):
        if kinds[qi] < 0.5:
            i = int(rng.integers(0, n))
            s = bytes(name_bytes[int(name_off[i]):int(name_off[i + 1])]).lower()
            k = s.rfind(b".")
            out.append(s[k + 1:])
        elif kinds[qi] < 0.9:
            out.append(vocab[int(rng.integers(0, len(vocab)))].encode("ascii"))
        else:
            # 3-4 letters from inside a vocabulary stem (selectivity around 1e-2; a window over the
            # constant "co.fanki.svc" prefix would match every name)
            w = vocab[int(rng.integers(0, len(vocab)))]
            ln = int(rng.integers(3, 5))
            st = int(rng.integers(0, len(w) - ln + 1))
            out.append(w[st:st + ln].encode("ascii"))
    return out


# ---- BASELINE config 4: the graph_query DSL mix ------------------------------------
C4_MIX = [("dependencies", 0.40), ("dependents", 0.20), ("overview", 0.15), ("khop", 0.10),
          ("resolve_name", 0.05), ("method", 0.05), ("entrypoints", 0.03), ("endpoints", 0.02)]
C4_OPCODES = [0, 1, 2, 3, 5, 6, 4, 7]   # DMG_OP_* of the kinds above, in C4_MIX order


def method_catalog(n_vertices, per_vertex=10):
    """10 methods per vertex held host-side: 8 lower-case bytes each, a 4-letter verb +
    4 hex digits of (7 v + j).  Returns (method_off u32[V+1], name_off u32[M+1], lower bytes u8[8 M])."""
    V = n_vertices
    verbs = np.frombuffer(b"findsavelistloadsendreadpushpullopenstop", dtype=np.uint8).reshape(10, 4)[:per_vertex]
    num = ((np.arange(V, dtype=np.int64)[:, None] * 7 + np.arange(per_vertex)[None, :]) & 0xFFFF)
    hexd = np.frombuffer(b"0123456789abcdef", dtype=np.uint8)
    digits = hexd[(num[:, :, None] >> np.array([12, 8, 4, 0])[None, None, :]) & 15]
    mbytes = np.concatenate([np.broadcast_to(verbs[None, :, :], (V, per_vertex, 4)), digits], axis=2).reshape(-1)
    return (np.arange(V + 1, dtype=np.uint32) * per_vertex, np.arange(per_vertex * V + 1, dtype=np.uint32) * 8,
            np.ascontiguousarray(mbytes))


def c4_mix_ops(n, names, mbytes, op_dtype, seed=SEED_C4 + 3, mode=0):
    """`n` tokenised graph_query operations in the SURVEY 8d config-4 mix over one graph of
    len(names) vertices: 40 % dependencies, 20 % dependents, 15 % overview, 10 % k-hop OUT k in
    {2,3,4}, 5 % simple-name resolution, 5 % ?check, 3 % entrypoints, 2 % endpoints.  Returns
    (ops structured array of op_dtype == dmg_query_op, text u8[], kind index per op into C4_MIX)."""
    V = len(names)
    r = np.random.Generator(np.random.PCG64(seed))
    u = r.random(n)
    kind = np.searchsorted(np.cumsum([w for _, w in C4_MIX]), u, side="right").clip(0, 7)
    opcode = np.array(C4_OPCODES, np.uint32)[kind]
    ops = np.zeros(n, op_dtype)
    ops["opcode"] = opcode
    ops["vertex"] = r.integers(0, V, n)
    ops["mode"] = mode
    ops["k"] = np.where(opcode == 3, r.integers(2, 5, n), 1)
    text = bytearray()
    for i in np.flatnonzero((opcode == 5) | (opcode == 6)):
        v = int(ops["vertex"][i])
        if opcode[i] == 5:
            t = names[v].rsplit(".", 1)[-1].lower().encode()
        elif r.random() < 0.8:
            j = int(r.integers(0, 10))
            t = bytes(mbytes[(v * 10 + j) * 8:(v * 10 + j + 1) * 8])
        else:
            t = b"nosuchmethod"
        ops["text_begin"][i] = len(text)
        text += t
        ops["text_end"][i] = len(text)
    return ops, np.frombuffer(bytes(text) + b"\0", dtype=np.uint8).copy(), kind
