"""CPU, world_size 2 over gloo: the N > 1 host logic of the partitioned closure (vertex-range
partition, edge ownership, per-level allgather of bitmap slices / contribution bitmaps, merge,
termination).  The per-level device kernels are replaced here by a numpy stand-in built on the
same CSR rules, so the orchestration that dmgraph/dist.py runs on GPUs is what is under test;
levels are checked against the CSR oracle."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class NumpyStepper:
    """Stand-in for dmg_closure_push_dev / pull_dev / merge_dev on CPU tensors."""

    def __init__(self, n, src, dst, v_lo, v_hi):
        self.n, self.v_lo, self.v_hi = n, v_lo, v_hi
        ok = (src < n) & (dst < n)
        key = np.unique(src[ok].astype(np.int64) * n + dst[ok])
        self.src, self.dst = key // n, key % n

    def _row_edges(self, direction, owned_side):
        a, b = (self.src, self.dst) if direction == 0 else (self.dst, self.src)     # a -> b along `direction`
        sel = (a >= self.v_lo) & (a < self.v_hi) if owned_side == "from" else (b >= self.v_lo) & (b < self.v_hi)
        return a[sel], b[sel]

    @staticmethod
    def _test(bits, v):
        return (bits[v >> 5] >> (v & 31)) & 1

    def push(self, direction, frontier, contrib):
        f = frontier.numpy().view(np.uint32)
        c = contrib.numpy().view(np.uint32)
        c[:] = 0
        a, b = self._row_edges(direction, "from")
        hit = self._test(f, a).astype(bool)
        np.bitwise_or.at(c, b[hit] >> 5, (1 << (b[hit] & 31)).astype(np.uint32))
        return int(hit.sum())

    def pull(self, direction, frontier, visited, next_slice):
        f = frontier.numpy().view(np.uint32)
        vis = visited.numpy().view(np.uint32)
        out = next_slice.numpy().view(np.uint32)
        out[:] = 0
        a, b = self._row_edges(direction, "to")
        hit = self._test(f, a).astype(bool) & ~self._test(vis, b).astype(bool)
        loc = b[hit] - self.v_lo
        np.bitwise_or.at(out, loc >> 5, (1 << (loc & 31)).astype(np.uint32))
        return int(a.size)

    def merge(self, lvl, gathered, n_gathered, stride, visited, nxt, level_slice):
        g = gathered.numpy().view(np.uint32)
        vis = visited.numpy().view(np.uint32)
        nx = nxt.numpy().view(np.uint32)
        words = vis.size
        x = np.zeros(words, np.uint32)
        for r in range(n_gathered):
            x |= g[r * stride: r * stride + words]
        fresh = x & ~vis
        nx[:] = fresh
        vis |= fresh
        v = np.nonzero(np.unpackbits(fresh.view(np.uint8), bitorder="little"))[0]
        own = v[(v >= self.v_lo) & (v < self.v_hi)] - self.v_lo
        level_slice.numpy()[own] = lvl
        return int(v.size)


def _worker(rank, world, port, n, src, dst, sources, direction, out_q):
    sys.path.insert(0, os.path.join(ROOT, "domain-mcp-server_b200"))
    from dmgraph import dist as ddist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    _wr, v_lo, v_hi = ddist.partition(n, rank, world)
    s_own, d_own = ddist.owned_edges(src, dst, v_lo, v_hi)
    stepper = NumpyStepper(n, s_own, d_own, v_lo, v_hi)
    pc = ddist.PartitionedClosure(n, rank, world, stepper, "cpu", dist)
    pc.PUSH_BELOW = 0.2                      # exercise both directions on a small graph
    level, reached = pc.run(sources, direction)
    out_q.put((rank, v_lo, v_hi, level.numpy().copy(), reached, dict(pc.stats)))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("direction", [0, 1])
def test_partitioned_closure_world2_gloo(direction):
    sys.path.insert(0, ROOT)
    from oracle import oracle as ora
    rng = np.random.default_rng(11 + direction)
    n, e = 1000, 2600
    src = rng.integers(0, n, e).astype(np.uint32)
    dst = rng.integers(0, n, e).astype(np.uint32)
    sources = rng.integers(0, n, 3).astype(np.uint32)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, src, dst, sources, direction, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want, reached, _te = ora.CsrOracle(n, src, dst).closure(direction, sources)
    level = np.full(n, 0xFFFFFFFF, np.uint32)
    for rank, v_lo, v_hi, lv, r, stats in got:
        assert r == reached
        level[v_lo:v_hi] = lv.view(np.uint32)[: v_hi - v_lo]
        assert stats["push_levels"] > 0 and stats["pull_levels"] > 0
    assert np.array_equal(level, want)


def test_partition_covers_all_vertices_and_edges():
    sys.path.insert(0, os.path.join(ROOT, "domain-mcp-server_b200"))
    from dmgraph import dist as ddist
    for n, world in [(1, 2), (31, 2), (64, 2), (1000, 8), (20_000_000, 8)]:
        spans = [ddist.partition(n, r, world)[1:] for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        for (a, b), (c, d) in zip(spans, spans[1:]):
            assert b == c and a % 32 == 0
    src = np.array([0, 40, 70, 5]); dst = np.array([40, 70, 0, 6])
    s0, d0 = ddist.owned_edges(src, dst, 0, 32)
    assert s0.tolist() == [0, 70, 5] and d0.tolist() == [40, 0, 6]
