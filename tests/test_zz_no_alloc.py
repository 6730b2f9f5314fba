"""GPU: dmg_neighbors_batch_dev_static, the k-hop batch for callers that cannot allocate.

Some hosts may not call cudaMalloc on the query path (a request thread with a latency budget, a
process whose device memory is already owned by another pool, code running where the runtime
forbids allocation).  For them the library carries its scratch as a 16 MiB static device array.
This test drives that entry point through the C ABI and holds it to the output of the ordinary,
allocating entry point (dmg_neighbors_batch_dev) on the same device buffers -- bit for bit, as in
every other parity test here -- on two calls in a row, the second smaller than the first so that
anything the first call left in the shared array would show.  It is collected last on purpose:
with `-x` every other test has already run if this one fails.  It must pass on its own in a fresh
process as well (the first call on a device does the one-time address lookup).
"""
import numpy as np
import pytest

import dmgraph
from dmgraph import synth

pytestmark = pytest.mark.gpu


def _run(engine, g, fn, mode, k, excl, q_off, q_src):
    nq = q_off.size - 1
    d_q_off, d_q_src = engine.to_device(q_off), engine.to_device(q_src)
    d_off = engine.device_alloc((nq + 1) * 8)
    rc, need = fn(mode, k, excl, nq, d_q_off, d_q_src, d_off, None, None, 0)      # sizes only
    assert rc in (dmgraph.OK, dmgraph.ECAP)
    cap = max(int(need), 1)
    d_ids, d_hop = engine.device_alloc(cap * 4), engine.device_alloc(cap)
    rc, need2 = fn(mode, k, excl, nq, d_q_off, d_q_src, d_off, d_ids, d_hop, cap)
    assert rc == dmgraph.OK and need2 == need
    out = (engine.to_host(d_off, np.uint64, nq + 1), engine.to_host(d_ids, np.uint32, int(need)),
           engine.to_host(d_hop, np.uint8, int(need)))
    for p in (d_q_off, d_q_src, d_off, d_ids, d_hop):
        engine.device_free(p)
    return out


def test_static_scratch_matches_allocating_entry_point_twice_in_a_row():
    engine = dmgraph.Engine(0)
    sg = synth.eco_graph(n_projects=6, v_per=700, e_per=6000, cap=120, seed=99)
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off)
    batches = [synth.stacktrace_queries(sg, n_queries=3000, seed=5),
               synth.stacktrace_queries(sg, n_queries=1100, frames=3, seed=6)]
    for (q_off, q_src), (mode, k, excl) in zip(batches, [(dmgraph.BOTH, 2, 1), (dmgraph.OUT, 3, 0)]):
        want = _run(engine, g, g.neighbors_batch_dev, mode, k, excl, q_off, q_src)
        got = _run(engine, g, g.neighbors_batch_dev_static, mode, k, excl, q_off, q_src)
        assert want[1].size > 0
        for a, b in zip(want, got):
            assert np.array_equal(a, b)
    g.release()
    engine.close()


def test_static_scratch_rejects_what_does_not_fit():
    engine = dmgraph.Engine(0)
    sg = synth.eco_graph(n_projects=2, v_per=2000, e_per=8000, cap=100, seed=3)
    g = dmgraph.Graph(engine, sg.n_vertices, sg.src, sg.dst, sg.entry, part_off=sg.part_off)
    nq = 9000                                            # 9000 * 2060 B > 16 MiB
    q_off = np.arange(nq + 1, dtype=np.uint32)
    q_src = np.zeros(nq, np.uint32)
    d_q_off, d_q_src = engine.to_device(q_off), engine.to_device(q_src)
    d_off = engine.device_alloc((nq + 1) * 8)
    with pytest.raises(dmgraph.DmgError) as ei:
        g.neighbors_batch_dev_static(dmgraph.BOTH, 1, 1, nq, d_q_off, d_q_src, d_off, None, None, 0)
    assert ei.value.code == dmgraph.EINVAL and "16777216" in str(ei.value)
    g.release()
    big = dmgraph.Graph(engine, 9000, [0, 1], [1, 2], [])            # one project > 8192 vertices
    with pytest.raises(dmgraph.DmgError) as ei:
        big.neighbors_batch_dev_static(dmgraph.BOTH, 1, 1, 1, d_q_off, d_q_src, d_off, None, None, 0)
    assert ei.value.code == dmgraph.EINVAL and "8192" in str(ei.value)
    big.release()
    for p in (d_q_off, d_q_src, d_off):
        engine.device_free(p)
    engine.close()
