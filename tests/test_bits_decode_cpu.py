"""CPU: dmg_bits_decode (host code of libdmgraph.so, no device work) against a plain-Python reading
of the bit-row format of include/dmgraph.h: bit v of plane b = bit b of hop(v)."""
import numpy as np

import dmgraph


def _rows(rng, nq, planes, words, max_wn):
    win = np.zeros((nq, 2), np.uint32)
    bits = np.zeros((nq, planes, words), np.uint32)
    want_ids, want_hop, counts = [], [], np.zeros(nq, np.uint32)
    for q in range(nq):
        wn = int(rng.integers(0, max_wn + 1))
        lo = int(rng.integers(0, 1 << 20))
        win[q] = (lo, wn)
        dens = rng.random()
        hop = np.where(rng.random(wn) < dens, rng.integers(1, 1 << planes, wn), 0)
        if rng.random() < 0.15:
            hop[:] = 0                                             # empty row
        for v in np.flatnonzero(hop):
            for b in range(planes):
                if (hop[v] >> b) & 1:
                    bits[q, b, v >> 5] |= np.uint32(1 << (v & 31))
        nz = np.flatnonzero(hop)
        counts[q] = nz.size
        want_ids.append(lo + nz)
        want_hop.append(hop[nz])
    return win, counts, bits, np.concatenate(want_ids).astype(np.uint32), np.concatenate(want_hop).astype(np.uint8)


def test_bits_decode_matches_the_format_definition():
    rng = np.random.default_rng(9)
    for planes, words, max_wn in [(1, 4, 128), (2, 64, 2000), (3, 8, 250), (2, 128, 4096)]:
        for threads in (1, 3, 0):
            win, counts, bits, wi, wh = _rows(rng, int(rng.integers(1, 200)), planes, words, max_wn)
            off, ids, hop = dmgraph.bits_decode(planes, words, win, counts, bits, threads=threads)
            assert np.array_equal(off, np.concatenate([[0], np.cumsum(counts, dtype=np.uint64)]))
            assert np.array_equal(ids, wi) and np.array_equal(hop, wh)


def test_bits_decode_capacity_answer():
    rng = np.random.default_rng(1)
    win, counts, bits, wi, _ = _rows(rng, 20, 2, 8, 256)
    if wi.size < 2:
        return
    off = np.zeros(21, np.uint64)
    small = (off, np.zeros(wi.size - 1, np.uint32), np.zeros(wi.size - 1, np.uint8))
    try:
        dmgraph.bits_decode(2, 8, win, counts, bits, out=small)
        assert False, "expected DMG_ECAP"
    except dmgraph.DmgError as e:
        assert e.code == dmgraph.ECAP


def test_bits_decode_streaming_path_on_a_large_result():
    """Results of more than 65,536 entries per thread go through the staged non-temporal stores: odd destination
    alignments (rows of odd lengths shift every thread's first byte), rows larger than the staging buffer, empty
    rows in between.  Expected arrays come from numpy on the same hop matrix."""
    rng = np.random.default_rng(4)
    for planes, wn, nq, dens in [(2, 2000, 900, 0.994), (3, 4096, 150, 1.0), (1, 333, 4000, 0.7), (2, 12000, 60, 0.9)]:
        words = (wn + 31) // 32 + 3
        hop = rng.integers(1, 1 << planes, (nq, wn)).astype(np.uint8)
        hop[rng.random((nq, wn)) >= dens] = 0
        hop[rng.integers(0, nq, nq // 10)] = 0                           # some empty rows
        lo = rng.integers(0, 1 << 20, nq).astype(np.uint32)
        win = np.stack([lo, np.full(nq, wn, np.uint32)], axis=1).astype(np.uint32)
        bits = np.zeros((nq, planes, words), np.uint32)
        pad = np.zeros((nq, words * 32), np.uint8)
        for b in range(planes):
            pad[:, :wn] = (hop >> b) & 1
            bits[:, b, :] = np.packbits(pad.reshape(nq, words, 32), axis=2, bitorder="little").view(np.uint32).reshape(nq, words)
        counts = (hop != 0).sum(1).astype(np.uint32)
        r, c = np.nonzero(hop)
        want_ids, want_hop = (lo[r] + c).astype(np.uint32), hop[r, c]
        for threads in (1, 3, 0):
            off, ids, hp = dmgraph.bits_decode(planes, words, win, counts, bits, threads=threads)
            assert np.array_equal(off, np.concatenate([[0], np.cumsum(counts, dtype=np.uint64)]))
            assert np.array_equal(ids, want_ids) and np.array_equal(hp, want_hop), (planes, wn, threads)
