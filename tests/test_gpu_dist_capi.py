"""dmg_dist_* on however many devices the box has (the round-end GPU tier has one; `tests/run_dist_capi.py` is
the same check at full size for 2 and 8 GPUs): a query batch sharded over the replicas and the vertex-range
partitioned closure in both transports -- the persistent peer-memory kernel and the per-level NCCL allgather --
against the oracle (PG:390-421 reachable set with depths; CCS:1563-1578 for the rows)."""
import os

import numpy as np
import pytest

import dmgraph
from dmgraph import synth
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dist():
    d = dmgraph.Dist(min(dmgraph.device_count(), 8))
    yield d
    d.close()


def test_sharded_batch_matches_oracle(dist):
    sg = synth.eco_graph(n_projects=6, v_per=700, e_per=6000, cap=80, seed=21, with_names=False)
    q_off, q_src = synth.stacktrace_queries(sg, n_queries=3001)          # does not divide by the device count
    dg = dmgraph.DistGraph(dist, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off)
    o = ora.CsrOracle(sg.n_vertices, sg.src, sg.dst, sg.entry, sg.part_off)
    for mode, k, excl in [(dmgraph.BOTH, 2, 1), (dmgraph.OUT, 3, 0), (dmgraph.IN, 1, 1)]:
        off, ids, hop = dg.neighbors_batch(mode, k, excl, q_off, q_src)
        woff, wids, whop, _ = o.khop_batch(mode, k, excl, q_off, q_src, threads=8)
        assert np.array_equal(off, woff) and np.array_equal(ids, wids) and np.array_equal(hop, whop), (mode, k, excl)
    dg.release()


@pytest.mark.parametrize("V,deg,n_src", [(200_003, 3, 5), (300_000, 12, 3000)])
def test_partitioned_closure_both_transports(dist, V, deg, n_src):
    rng = np.random.Generator(np.random.PCG64(V))
    src = rng.integers(0, V, V * deg).astype(np.uint32)
    dst = rng.integers(0, V, V * deg).astype(np.uint32)
    srcs = rng.integers(0, V + 10, n_src).astype(np.uint32)               # a few unknown ids (ignored, PG:194-196)
    pg = dmgraph.DistGraph(dist, V, src, dst, srcs[srcs < V], partitioned=True)
    o = ora.CsrOracle(V, src, dst, srcs[srcs < V])
    try:
        for direction in (dmgraph.OUT, dmgraph.IN):
            wlvl, wreached, _ = o.closure(direction, srcs)
            for env, transport in ((None, 1), ("1", 0)):
                if env:
                    os.environ["DMG_DIST_NCCL"] = env
                else:
                    os.environ.pop("DMG_DIST_NCCL", None)
                for _ in range(2):                                        # the second call reuses the rank state
                    lvl, reached, st = pg.closure(direction, srcs)
                    assert reached == wreached and np.array_equal(lvl, wlvl), (direction, env)
                assert st["transport"] == transport and st["levels"] == int(wlvl[wlvl != 0xFFFFFFFF].max()) + 1
    finally:
        os.environ.pop("DMG_DIST_NCCL", None)
        pg.release()


@pytest.mark.parametrize("V,deg,n_entry", [(100_001, 4, 7), (250_000, 10, 1500), (40_000, 2, 0)])
def test_partitioned_analysis_order_is_exact(dist, V, deg, n_entry):
    """dmg_dist_analysis_order == the oracle's FIFO order (PG:390-421), orphans in nodes order included."""
    rng = np.random.Generator(np.random.PCG64(V + 5))
    src = rng.integers(0, V, V * deg).astype(np.uint32)
    dst = rng.integers(0, V, V * deg).astype(np.uint32)
    entry = rng.permutation(V)[:n_entry].astype(np.uint32)                 # insertion order is not id order
    pg = dmgraph.DistGraph(dist, V, src, dst, entry, partitioned=True)
    o = ora.CsrOracle(V, src, dst, entry)
    want = o.analysis_order()
    for _ in range(2):
        got = pg.analysis_order()
        assert np.array_equal(got, want)
    pg.release()


def test_sharded_query_batch_equals_one_device(dist):
    """dmg_dist_query_batch (config 4: DSL batches sharded over the replicas) gives the offsets, ids and aux
    bytes of ONE dmg_query_batch call on one device, which tests/test_gpu_fullsize.py holds to the oracle
    opcode by opcode; a sample of rows is checked against the oracle here as well."""
    from oracle import checks
    sg = synth.big_graph(30_000, 400_000, cap=200, entry_frac=0.01, seed=91, with_names=True)
    V = sg.n_vertices
    off, arena = synth.pack_names(sg.names)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    mo, no, mbytes = synth.method_catalog(V)
    flags = np.repeat((np.random.Generator(np.random.PCG64(4)).random(V) < 0.02).astype(np.uint8), 10)
    ops, text, kind = synth.c4_mix_ops(3000, sg.names, mbytes, dmgraph.Graph.QUERY_OP_DTYPE)
    dg = dmgraph.DistGraph(dist, V, sg.src, sg.dst, sg.entry, name_off=off, name_bytes=arena, lower_bytes=low)
    for r in range(dist.n):
        part = dg.part(r)
        part.set_methods_packed(mo, no, mbytes)
        part.set_method_flags(flags)
    need = dg.query_batch_packed(ops, text)
    got = (np.zeros(len(ops) + 1, np.uint64), np.zeros(max(need, 1), np.uint32), np.zeros(max(need, 1), np.uint8))
    assert dg.query_batch_packed(ops, text, out=got) == need
    one = dg.part(0)
    want = (np.zeros(len(ops) + 1, np.uint64), np.zeros(max(need, 1), np.uint32), np.zeros(max(need, 1), np.uint8))
    assert one.query_batch_packed(ops, text, out=want) == need
    assert all(np.array_equal(a, b) for a, b in zip(got, want))
    o = ora.CsrOracle(V, sg.src, sg.dst, sg.entry)
    rows = np.sort(np.random.Generator(np.random.PCG64(8)).choice(len(ops), 400, replace=False))
    assert checks.check_query_mix(o, ops, text, got[0], got[1], got[2], off, low, mbytes,
                                  np.flatnonzero(flags).astype(np.uint32), rows=rows) == rows.size
    dg.release()
