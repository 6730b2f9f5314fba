"""Oracle parity at the sizes BASELINE.json states: the CUDA path through
the C ABI against the CPU restatements on the SAME full-size inputs bench.py times -- not a slice,
not the engine against itself.  Bit-exact ids, hops, orders.

  C2  the whole 65,536-query stack-trace batch over 100 k classes / 2 M edges (BOTH, k = 2)
  C3  2 M catalog names x all 16,384 queries, ordered hits
  C4  one 1 M / 30 M graph: dependents incl. the hub rows, directed k-hop k = 2, 3, and a
      dmg_query_batch mix in which every opcode's rows are checked
  C5  closure levels + exact analysisOrder on 4 M vertices / 100 M drawn edges
"""
import numpy as np
import pytest

import dmgraph
from dmgraph import synth
from oracle import checks
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    e = dmgraph.Engine(0)
    yield e
    e.close()


def test_config2_whole_batch_against_oracle(engine):
    sg = synth.eco_graph(with_names=False)
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off)
    o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry, sg.part_off)
    q_off, q_src = synth.stacktrace_queries(sg, n_queries=65536)
    off, ids, hop = g.neighbors_batch(dmgraph.BOTH, 2, 1, q_off, q_src)
    assert checks.check_khop_rows(o, 2, 2, 1, q_off, q_src, off, ids, hop) == 65536
    # k = 1 is the reference's own resolveNeighbors (CCS:1558-1578): the whole batch as well
    off, ids, hop = g.neighbors_batch(dmgraph.BOTH, 1, 1, q_off, q_src)
    assert checks.check_khop_rows(o, 2, 1, 1, q_off, q_src, off, ids, hop) == 65536
    g.release()


def test_config3_two_million_names_all_queries(engine):
    n = 2_000_000
    off, arena = synth.catalog_names(n)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    queries = synth.search_queries(off, arena, 16384)
    g = dmgraph.Graph(engine, n, [], [], [], name_off=off, name_bytes=arena, lower_bytes=low)
    o, ids = g.search_batch(dmgraph.SEARCH_CONTAINS_ALL, queries)
    wo, wids = ora.search_batch_ac(off, low, queries, threads=ora.hardware_threads())
    assert int(wo[-1]) > 10 * n                      # ~25 hits per name on this query mix
    assert np.array_equal(o, wo) and np.array_equal(ids, wids)
    # the restatement proper (one contains() per query and name) on a sample of the same batch
    sample = list(range(0, 16384, 512))
    so, sids = ora.search_batch(off, low, [queries[i] for i in sample], 0, threads=ora.hardware_threads())
    for j, i in enumerate(sample):
        assert np.array_equal(ids[int(o[i]):int(o[i + 1])], sids[int(so[j]):int(so[j + 1])])
    g.release()


@pytest.fixture(scope="module")
def c4(engine):
    sg = synth.big_graph(1_000_000, 30_000_000, cap=500, entry_frac=0.01, seed=synth.SEED_C4, with_names=True)
    off, arena = synth.pack_names(sg.names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry, name_off=off, name_bytes=arena, lower_bytes=low)
    o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry)
    yield sg, g, o, off, low
    g.release()


def test_config4_dependents_with_hub_rows(engine, c4):
    sg, g, o, _off, _low = c4
    rng = np.random.Generator(np.random.PCG64(synth.SEED_C4 + 1))
    indeg = np.bincount(sg.dst, minlength=sg.n_vertices)
    hubs = np.argsort(indeg)[-16:]
    assert indeg[hubs].max() > 100_000               # Zipf head: rows of several hundred thousand dependents
    ids = np.concatenate([rng.integers(0, sg.n_vertices, 262_144 - 16), hubs]).astype(np.uint32)
    for d in (dmgraph.OUT, dmgraph.IN):
        goff, gids = g.adjacency_batch(d, ids)
        woff, wids = o.adjacency_batch(d, ids)
        assert np.array_equal(goff, woff) and np.array_equal(gids, wids)


@pytest.mark.parametrize("k", [1, 2, 3])
def test_config4_directed_khop(engine, c4, k):
    sg, g, o, _off, _low = c4
    rng = np.random.Generator(np.random.PCG64(synth.SEED_C4 + 10 + k))
    nq = 4096
    src = rng.integers(0, sg.n_vertices, nq).astype(np.uint32)
    q_off = np.arange(nq + 1, dtype=np.uint32)
    off, ids, hop = g.neighbors_batch(dmgraph.OUT, k, 0, q_off, src)
    assert checks.check_khop_rows(o, 0, k, 0, q_off, src, off, ids, hop) == nq
    if k == 2:                                        # dependents direction and resolve() on the same batch
        for mode in (dmgraph.IN, dmgraph.BOTH):
            off, ids, hop = g.neighbors_batch(mode, k, 1, q_off[:513], src[:512])
            assert checks.check_khop_rows(o, mode, k, 1, q_off[:513], src[:512], off, ids, hop) == 512


def test_config4_query_mix_every_opcode(engine, c4):
    sg, g, o, off, low = c4
    V = sg.n_vertices
    mo, no, mbytes = synth.method_catalog(V)
    g.set_methods_packed(mo, no, mbytes)
    rng = np.random.Generator(np.random.PCG64(synth.SEED_C4 + 2))
    ctrl = rng.random(V) < 0.01
    flags = np.repeat(ctrl.astype(np.uint8), 10)
    g.set_method_flags(flags)
    endpoint_methods = np.flatnonzero(flags).astype(np.uint32)
    n = 8192
    ops, text, kind = synth.c4_mix_ops(n, sg.names, mbytes, dmgraph.Graph.QUERY_OP_DTYPE)
    assert set(kind.tolist()) == set(range(8))        # every opcode of the mix is present
    need = g.query_batch_packed(ops, text)
    o_off = np.zeros(n + 1, np.uint64)
    ids, aux = np.zeros(need, np.uint32), np.zeros(need, np.uint8)
    assert g.query_batch_packed(ops, text, out=(o_off, ids, aux)) == need
    assert checks.check_query_mix(o, ops, text, o_off, ids, aux, off, low, mbytes, endpoint_methods) == n


def test_config5_closure_and_exact_order_4m(engine):
    sg = synth.big_graph_fast(4_000_000, 100_000_000, cap=2000, entry_frac=0.01, seed=synth.SEED_C5)
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry)
    o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry)
    assert g.info["n_edges"] == o.n_edges and g.info["duplicate_edges"] > 0
    lvl, reached = g.closure(dmgraph.OUT, sg.entry)
    wlvl, wreached, wte = o.closure(0, sg.entry)
    assert reached == wreached and np.array_equal(lvl, wlvl)
    assert engine.last_stats()["edges_traversed"] == wte
    assert np.array_equal(g.analysis_order(), o.analysis_order())          # PG:390-421, exact FIFO order
    rng = np.random.Generator(np.random.PCG64(synth.SEED_C5 + 1))
    for s in rng.integers(0, sg.n_vertices, 4):                             # single-source closures, both directions
        for d in (dmgraph.OUT, dmgraph.IN):
            lvl, reached = g.closure(d, [int(s)])
            wlvl, wreached, _ = o.closure(d, [int(s)])
            assert reached == wreached and np.array_equal(lvl, wlvl)
    g.release()
