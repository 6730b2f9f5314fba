"""k-hop on projects beyond the shared-memory classes (> 8192 vertices): khop1_warp_kernel (k = 1, one warp
per query, hub rows and long source lists handed to the CTA kernel), khop_large_kernel (claim bitmap in
shared memory) and khop_global_kernel (claim bitmap in global memory; forced here with DMG_NO_KHOP_LARGE
because only projects beyond ~1.4 M vertices reach it otherwise).  Every row is compared with the oracle
(semantics: CCS:1563-1578 iterated k times)."""
import os

import numpy as np
import pytest

import dmgraph
from oracle import oracle as ora

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    e = dmgraph.Engine(0)
    yield e
    e.close()


def _graph(n_parts, v_per, deg, seed):
    """Projects of v_per vertices; vertex 7 of each project is a hub both ways (rows of > 1024 entries)."""
    rng = np.random.default_rng(seed)
    src, dst = [], []
    for p in range(n_parts):
        b = p * v_per
        s = rng.integers(0, v_per, v_per * deg)
        d = rng.integers(0, v_per, v_per * deg)
        hub_in = rng.choice(v_per, 1500, replace=False)
        hub_out = rng.choice(v_per, 1300, replace=False)
        src += [b + s, b + hub_in, np.full(1300, b + 7)]
        dst += [b + d, np.full(1500, b + 7), b + hub_out]
    src = np.concatenate(src).astype(np.uint32)
    dst = np.concatenate(dst).astype(np.uint32)
    part_off = (np.arange(n_parts + 1) * v_per).astype(np.uint32)
    return src, dst, part_off


def _queries(rng, n_parts, v_per, n):
    """Ragged source lists: single sources, the hub, repeated sources, lists longer than a warp, sources of
    another project after the first (they match nothing: quirk 6), ids beyond V, empty lists."""
    V = n_parts * v_per
    rows = []
    for i in range(n):
        p = int(rng.integers(0, n_parts))
        b = p * v_per
        kind = i % 8
        if kind == 0:
            r = [b + 7]
        elif kind == 1:
            r = list(b + rng.integers(0, v_per, 3)) + [b + 7]
        elif kind == 2:
            x = int(b + rng.integers(0, v_per))
            r = [x, x, int(b + rng.integers(0, v_per)), x]
        elif kind == 3:
            r = list(b + rng.integers(0, v_per, 40))
        elif kind == 4:
            other = ((p + 1) % n_parts) * v_per
            r = [int(b + rng.integers(0, v_per)), int(other + rng.integers(0, v_per)), V + 5]
        elif kind == 5:
            r = []
        else:
            r = list(b + rng.integers(0, v_per, int(rng.integers(1, 6))))
        rows.append(np.asarray(r, np.uint32))
    q_off = np.zeros(n + 1, np.uint32)
    np.cumsum([r.size for r in rows], out=q_off[1:])
    return q_off, (np.concatenate(rows) if rows else np.zeros(0, np.uint32))


@pytest.mark.parametrize("n_parts,v_per,deg", [(1, 9001, 6), (3, 20000, 5), (1, 150000, 8)])
@pytest.mark.parametrize("force", ["DMG_FORCE_KHOP_LARGE", "DMG_NO_KHOP_LARGE", None])
def test_large_project_khop_matches_oracle(engine, n_parts, v_per, deg, force):
    src, dst, part_off = _graph(n_parts, v_per, deg, seed=v_per)
    V = n_parts * v_per
    entry = np.array([0], np.uint32)
    g = dmgraph.Graph(engine, V, src, dst, entry, part_off=part_off)
    o = ora.CsrOracle(V, src, dst, entry, part_off)
    q_off, q_src = _queries(np.random.default_rng(v_per + 1), n_parts, v_per, 400)
    if force:
        os.environ[force] = "1"
    try:
        for mode in (dmgraph.OUT, dmgraph.IN, dmgraph.BOTH):
            # k = 6 only one way: deeper than the level boundaries the emit keeps in registers
            for k, excl in [(1, 0), (1, 1), (2, 1), (3, 0)] + ([(6, 1)] if mode == dmgraph.OUT else []):
                off, ids, hop = g.neighbors_batch(mode, k, excl, q_off, q_src)
                st = engine.last_stats()
                woff, wids, whop, wte = o.khop_batch(mode, k, excl, q_off, q_src, threads=8)
                assert np.array_equal(off, woff), (mode, k, excl)
                assert np.array_equal(ids, wids) and np.array_equal(hop, whop), (mode, k, excl)
                assert st["edges_traversed"] == wte, (mode, k, excl)
    finally:
        if force:
            os.environ.pop(force, None)
    g.release()


def test_large_project_capacity_and_sizing(engine):
    """A call without room answers DMG_ECAP with `needed`; the retry with that capacity gives the rows."""
    src, dst, part_off = _graph(1, 30000, 6, seed=3)
    g = dmgraph.Graph(engine, 30000, src, dst, np.array([1], np.uint32))
    o = ora.CsrOracle(30000, src, dst, np.array([1], np.uint32))
    q_off, q_src = _queries(np.random.default_rng(9), 1, 30000, 128)
    for k in (1, 3):
        off, ids, hop = g.neighbors_batch(dmgraph.BOTH, k, 0, q_off, q_src)     # sizes, then fetches
        woff, wids, whop, _ = o.khop_batch(2, k, 0, q_off, q_src, threads=8)
        assert np.array_equal(off, woff) and np.array_equal(ids, wids) and np.array_equal(hop, whop)
    g.release()
