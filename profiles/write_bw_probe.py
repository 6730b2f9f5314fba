#!/usr/bin/env python3
"""How fast can this B200 take a write-only stream?  khop_emit writes ~600 MB of packed rows per
config-2 step and reads ~140 MB; three different kernel bodies all took 0.24 ms, so this probe
times plain torch fills/copies of the same size with the L2 in the same states bench.py leaves it:
dirty (after a write flush) and clean (after a read sweep)."""
import json
import torch

dev = torch.device("cuda:0")
n = 600 << 20
out = torch.empty(n, dtype=torch.uint8, device=dev)
src = torch.empty(n, dtype=torch.uint8, device=dev)
big = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timed(prep, fn, reps=10):
    ts = []
    for _ in range(reps):
        prep()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


dirty = lambda: big.fill_(1)
clean = lambda: big.sum()
res = {}
for name, prep in (("dirty_l2", dirty), ("clean_l2", clean)):
    ms = timed(prep, lambda: out.zero_())
    res["fill_600MB_" + name] = {"ms": ms, "GBs": n / ms / 1e6}
    ms = timed(prep, lambda: out[: n // 2].copy_(src[: n // 2]))
    res["copy_300MB_to_300MB_" + name] = {"ms": ms, "GBs": n / ms / 1e6}
print(json.dumps(res))
