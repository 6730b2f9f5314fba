"""The Aho-Corasick form of the search oracle (used for full-size config 3 parity) is held to the
restatement of the reference's loop -- one contains() per (query, identifier), CCS:1511-1526 --
on random inputs: overlapping and nested patterns, repeated queries, empty queries, ranges."""
import numpy as np

from oracle import oracle as ora


def _arena(rng, n, alphabet, max_len):
    names = [bytes(rng.choice(alphabet, int(rng.integers(0, max_len + 1))).tolist()) for _ in range(n)]
    off = np.zeros(n + 1, np.uint64)
    np.cumsum([len(x) for x in names], out=off[1:])
    return names, off, np.frombuffer(b"".join(names) + b"\0", dtype=np.uint8)[:-1].copy()


def test_ac_equals_contains_loop_on_random_inputs():
    rng = np.random.default_rng(11)
    for trial in range(30):
        alphabet = np.frombuffer(b"ab.c"[: int(rng.integers(1, 5))], dtype=np.uint8)
        names, off, low = _arena(rng, int(rng.integers(1, 300)), alphabet, 24)
        queries = []
        for _ in range(int(rng.integers(1, 80))):
            r = rng.random()
            if r < 0.1:
                queries.append(b"")
            elif r < 0.6 and names:
                nm = names[int(rng.integers(0, len(names)))]
                a = int(rng.integers(0, len(nm) + 1))
                queries.append(nm[a:a + int(rng.integers(0, 9))])
            else:
                queries.append(bytes(rng.choice(alphabet, int(rng.integers(1, 6))).tolist()))
        queries += queries[:3]                                        # repeated query texts
        lo = int(rng.integers(0, len(names)))
        hi = int(rng.integers(lo, len(names) + 1))
        for v_lo, v_hi in [(0, len(names)), (lo, hi), (lo, lo)]:
            a = ora.search_batch(off, low, queries, 0, v_lo, v_hi, threads=2)
            b = ora.search_batch_ac(off, low, queries, v_lo, v_hi, threads=int(rng.integers(1, 5)))
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_ac_on_config3_shaped_names():
    from dmgraph import synth
    off, arena = synth.catalog_names(20000)
    low = np.frombuffer(bytes(arena).lower(), dtype=np.uint8).copy()
    qs = synth.search_queries(off, arena, 400)
    a = ora.search_batch(off, low, qs, 0, threads=4)
    b = ora.search_batch_ac(off, low, qs, threads=3)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[1].size > 0
