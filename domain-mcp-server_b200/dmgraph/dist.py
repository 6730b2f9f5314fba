"""Vertex-range partitioned closure over the GPUs of one box (BASELINE config 5).

One process per GPU (torchrun).  torch.distributed is the plumbing only: it owns the bitmap
tensors and runs the per-level allgather (NCCL over NVLink; gloo in the CPU tests).  Every level's
work -- push / pull expansion, merge, level stamping -- is the C ABI's
dmg_closure_push_dev / dmg_closure_pull_dev / dmg_closure_merge_dev on this rank's device.

    rank r owns vertex words [r*Wr, (r+1)*Wr)  (Wr = ceil(ceil(V/32) / world), 32 vertices a word)
    frontier / visited / next : world*Wr words, replicated on every rank
    per level:  push  -> all_gather(full contribution bitmaps) -> merge(OR of `world` bitmaps)
                pull  -> all_gather(own slices)                -> merge(1 bitmap)
"""
import numpy as np

OUT, IN = 0, 1


def partition(n_vertices, rank, world):
    """(words_per_rank, v_lo, v_hi) of `rank`; ranges start on multiples of 32."""
    words = (n_vertices + 31) // 32
    wr = max(1, (words + world - 1) // world)
    v_lo = min(n_vertices, rank * wr * 32)
    v_hi = min(n_vertices, (rank + 1) * wr * 32)
    return wr, v_lo, v_hi


def owned_edges(src, dst, v_lo, v_hi):
    """Edges a rank uploads: every edge with an endpoint in its range, original order kept, so
    the out rows and the in rows of its own vertices are complete."""
    src = np.asarray(src)
    dst = np.asarray(dst)
    keep = ((src >= v_lo) & (src < v_hi)) | ((dst >= v_lo) & (dst < v_hi))
    return src[keep], dst[keep]


class GpuStepper:
    """Level primitives on this rank's GPU through the C ABI; bitmaps are torch CUDA int32 tensors."""

    def __init__(self, graph, v_lo, v_hi):
        self.g, self.v_lo, self.v_hi = graph, v_lo, v_hi

    def push(self, direction, frontier, contrib):
        import torch
        torch.cuda.current_stream().synchronize()
        return self.g.closure_push_dev(direction, self.v_lo, self.v_hi, frontier.data_ptr(), contrib.data_ptr())

    def pull(self, direction, frontier, visited, next_slice):
        import torch
        torch.cuda.current_stream().synchronize()
        return self.g.closure_pull_dev(direction, self.v_lo, self.v_hi, frontier.data_ptr(), visited.data_ptr(),
                                       next_slice.data_ptr())

    def merge(self, lvl, gathered, n_gathered, stride, visited, nxt, level_slice):
        import torch
        torch.cuda.current_stream().synchronize()
        return self.g.closure_merge_dev(self.v_lo, self.v_hi, lvl, gathered.data_ptr(), n_gathered, stride,
                                        visited.data_ptr(), nxt.data_ptr(), level_slice.data_ptr())


class GpuStreamStepper:
    """The same primitives enqueued on torch's current CUDA stream (the stream NCCL runs on): no
    synchronisation inside a level; PartitionedClosure reads |next| once per level."""

    asynchronous = True

    def __init__(self, graph, v_lo, v_hi):
        import ctypes as C
        import torch
        self.g, self.v_lo, self.v_hi, self.C, self.torch = graph, v_lo, v_hi, C, torch
        self.edges = torch.zeros(1, dtype=torch.int64, device="cuda")
        self.n_new = torch.zeros(1, dtype=torch.int64, device="cuda")

    def _stream(self):
        return self.C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def _p(self, t):
        return self.C.c_void_p(t.data_ptr())

    def push(self, direction, frontier, contrib):
        from dmgraph import _check
        _check(self.g.L.dmg_closure_push_async(self.g.h, direction, self.v_lo, self.v_hi, self._p(frontier),
                                               self._p(contrib), self._p(self.edges), self._stream()))
        return 0

    def pull(self, direction, frontier, visited, next_slice):
        from dmgraph import _check
        _check(self.g.L.dmg_closure_pull_async(self.g.h, direction, self.v_lo, self.v_hi, self._p(frontier),
                                               self._p(visited), self._p(next_slice), self._p(self.edges),
                                               self._stream()))
        return 0

    def merge(self, lvl, gathered, n_gathered, stride, visited, nxt, level_slice):
        from dmgraph import _check
        _check(self.g.L.dmg_closure_merge_async(self.g.h, self.v_lo, self.v_hi, lvl, self._p(gathered), n_gathered,
                                                stride, self._p(visited), self._p(nxt), self._p(level_slice),
                                                self._p(self.n_new), self._stream()))
        return int(self.n_new.item())          # the one synchronisation of the level

    def take_edges(self):
        e = int(self.edges.item())
        self.edges.zero_()
        return e


class PartitionedClosure:
    """BFS levels from a source set over `direction`, the graph split by vertex range over ranks."""

    PUSH_BELOW = 1.0 / 32     # push while the frontier is below this fraction of V

    def __init__(self, n_vertices, rank, world, stepper, device, dist=None):
        import torch
        self.torch, self.dist = torch, dist
        self.V, self.rank, self.world, self.stepper = n_vertices, rank, world, stepper
        self.wr, self.v_lo, self.v_hi = partition(n_vertices, rank, world)
        self.words = self.wr * world                    # padded bitmap width
        z = lambda n: torch.zeros(max(n, 1), dtype=torch.int32, device=device)
        self.frontier, self.visited, self.next = z(self.words), z(self.words), z(self.words)
        self.contrib = z(self.words)
        self.gather_push = z(self.words * world)
        self.slice = z(self.wr)
        self.level = torch.full((max(self.v_hi - self.v_lo, 1),), -1, dtype=torch.int32, device=device)
        self.stats = {"levels": 0, "push_levels": 0, "pull_levels": 0, "edges": 0, "exchanged_bytes": 0}

    def _allgather(self, out, inp):
        if self.dist is None or self.world == 1:
            out.copy_(inp.repeat(out.numel() // inp.numel())[:out.numel()] if out.numel() != inp.numel() else inp)
        else:
            self.dist.all_gather_into_tensor(out, inp)
        self.stats["exchanged_bytes"] += out.numel() * 4

    def run(self, sources, direction=OUT):
        torch = self.torch
        # source bitmap, built where the bitmaps live (no host-side bit twiddling on V-sized arrays)
        dev = self.frontier.device
        s = torch.as_tensor(np.asarray(sources, dtype=np.int64), device=dev)
        s = torch.unique(s[(s >= 0) & (s < self.V)])
        mask = torch.zeros(self.words * 32, dtype=torch.int64, device=dev)
        mask[s] = 1
        packed = (mask.view(-1, 32) << torch.arange(32, dtype=torch.int64, device=dev)).sum(dim=1)
        packed = torch.where(packed >= 2 ** 31, packed - 2 ** 32, packed).to(torch.int32)
        self.frontier.copy_(packed)
        self.visited.copy_(packed)
        self.level.fill_(-1)
        own = s[(s >= self.v_lo) & (s < self.v_hi)] - self.v_lo
        if own.numel():
            self.level[own] = 0
        s = s.cpu().numpy()
        import time
        n_front, reached, lvl = int(s.size), int(s.size), 1
        self.stats["level_ms"] = []
        while n_front > 0:
            t_lvl = time.perf_counter()
            mode = "push" if n_front < self.PUSH_BELOW * self.V else "pull"
            if n_front < self.PUSH_BELOW * self.V:
                self.stats["edges"] += self.stepper.push(direction, self.frontier, self.contrib)
                self._allgather(self.gather_push, self.contrib)
                n_new = self.stepper.merge(lvl, self.gather_push, self.world, self.words, self.visited, self.next,
                                           self.level)
                self.stats["push_levels"] += 1
            else:
                self.stats["edges"] += self.stepper.pull(direction, self.frontier, self.visited, self.slice)
                self._allgather(self.contrib, self.slice)      # contrib doubles as the gathered next bitmap
                n_new = self.stepper.merge(lvl, self.contrib, 1, self.words, self.visited, self.next, self.level)
                self.stats["pull_levels"] += 1
            self.frontier, self.next = self.next, self.frontier
            self.stats["level_ms"].append((mode, n_front, round(1e3 * (time.perf_counter() - t_lvl), 3)))
            n_front = int(n_new)
            reached += n_front
            lvl += 1
        self.stats["levels"] = lvl - 1
        if hasattr(self.stepper, "take_edges"):
            self.stats["edges"] += self.stepper.take_edges()
        return self.level[: self.v_hi - self.v_lo], reached
