"""CPU model of the arithmetic of k_octscan (csrc/level.cuh): the packed counters must give exactly the plain segmented
eight-counter scan that OctScanOp defines (rank of a loser inside its (cell, octant) bucket, bucket totals at the last
arrival of a cell).  The kernel itself is compared with the oracle on the GPU (tests/test_gpu_parity.py); this test pins
the packing: 8-bit counters per thread, two 16-bit fields per register across the tile with the "cell starts here" flag
riding in bit 31 of the fourth register, 13-bit field extraction, tile prefix added only where no cell started before.
"""
import numpy as np
import pytest

TILE_THREADS, ITEMS = 256, 16
TILE = TILE_THREADS * ITEMS
HEAD = 0x80000000
M32 = 0xFFFFFFFF


def plain_scan(octs, first):
    """rank of every emitting arrival inside its (cell, octant) bucket + inclusive counters after every arrival"""
    run = [0] * 8
    ranks, incl = [], []
    for o, f in zip(octs, first):
        if f:
            run = [0] * 8
        r = 0
        if o < 8:
            r = run[o]
            run[o] += 1
        ranks.append(r)
        incl.append(list(run))
    return ranks, incl


def byte_step(code_oct, code_first, lo, hi, started):
    """oct_bytes_step: returns (count of the octant before this arrival, lo, hi, started)"""
    if code_first:
        lo, hi, started = 0, 0, 1
    sel = code_oct & 3
    w = hi if (code_oct & 4) else lo
    r = (w >> (8 * sel)) & 0xFF          # __byte_perm(w, 0, 0x4440 | sel)
    one = 1 << (8 * sel)
    if code_oct < 4:
        lo = (lo + one) & M32
    if code_oct & 4:                     # octant 8 (no emission) has neither bit: counts nothing
        hi = (hi + one) & M32
    return r, lo, hi, started


def spread(b):   # __byte_perm(b, 0, 0x4140) / 0x4342: two bytes -> two 16-bit fields
    return (b & 0xFF) | (((b >> 8) & 0xFF) << 16), ((b >> 16) & 0xFF) | (((b >> 24) & 0xFF) << 16)


def combine(a, b):   # octw_combine
    if b[3] & HEAD:
        return list(b)
    return [(x + y) & M32 for x, y in zip(a, b)]


def field(w, q):     # octw_field
    return (w[q >> 1] >> (16 * (q & 1))) & 0x1FFF


def kernel_model(octs, first, last, prefix):
    """One tile of k_octscan with the tile prefix `prefix` (8 counters of the tiles before, since the last cell start).
    Returns ranks, {position of a cell's last arrival: bucket totals}, the tile's aggregate (8 counters, head)."""
    n = len(octs)
    assert n <= TILE
    octs = list(octs) + [8] * (TILE - n)
    first = list(first) + [0] * (TILE - n)
    last = list(last) + [0] * (TILE - n)
    # per thread: totals since the last cell start inside the thread
    tw = []
    for t in range(TILE_THREADS):
        lo = hi = started = 0
        for k in range(ITEMS):
            i = t * ITEMS + k
            _, lo, hi, started = byte_step(octs[i], first[i], lo, hi, started)
        a0, a1 = spread(lo)
        a2, a3 = spread(hi)
        tw.append([a0, a1, a2, a3 | (HEAD if started else 0)])
    # exclusive segmented scan over the threads (the kernel does warp shuffles + a cross-warp fold: any order of an
    # associative operator gives the same result)
    excl, acc = [], [0, 0, 0, 0]
    for t in range(TILE_THREADS):
        excl.append(list(acc))
        acc = combine(acc, tw[t])
    total = acc
    ranks, totals = [], {}
    for t in range(TILE_THREADS):
        tcut = bool(excl[t][3] & HEAD)
        start = [field(excl[t], q) for q in range(8)]       # in-tile start values (s_run column)
        lo = hi = started = 0
        for k in range(ITEMS):
            i = t * ITEMS + k
            o = octs[i]
            local, lo, hi, started = byte_step(o, first[i], lo, hi, started)
            r = local if started else start[o if o < 8 else 0] + local
            nofix = bool(started) or tcut or o >= 8
            word = (r if o < 8 else 0) | (o << 16) | (int(nofix) << 20)
            assert (r if o < 8 else 0) < (1 << 16)
            # on the way out
            rr = (word & 0xFFFF) + (0 if (word >> 20) & 1 else prefix[(word >> 16) & 7])
            if i < n:
                ranks.append(rr)
                if last[i]:   # oct_cell_totals
                    tot = []
                    for q in range(8):
                        v = ((lo if q < 4 else hi) >> (8 * (q & 3))) & 0xFF
                        if not started:
                            v += start[q] + (0 if tcut else prefix[q])
                        tot.append(v)
                    totals[i] = tot
    return ranks, totals, [field(total, q) for q in range(8)], bool(total[3] & HEAD)


def make_case(rng, n, p_first, weights):
    octs = rng.choice(9, size=n, p=weights).tolist()
    first = (rng.random(n) < p_first).astype(int).tolist()
    last = [0] * n
    for i in range(n - 1):
        if first[i + 1]:
            last[i] = 1
    return octs, first, last


@pytest.mark.parametrize("seed", range(6))
def test_tile_equals_plain_segmented_scan(seed):
    rng = np.random.default_rng(seed)
    n = [TILE, TILE, 4001, 17, 1, 3000][seed]
    w = rng.random(9) + 0.05
    w /= w.sum()
    octs, first, last = make_case(rng, n, [0.0, 0.002, 0.01, 0.2, 0.0, 0.0005][seed], w)
    prefix = [int(x) for x in rng.integers(0, 1 << 20, 8)]
    ranks, totals, agg, head = kernel_model(octs, first, last, prefix)
    # plain scan that starts from the prefix
    run = list(prefix)
    for i in range(n):
        if first[i]:
            run = [0] * 8
        o = octs[i]
        want = 0
        if o < 8:
            want = run[o]
            run[o] += 1
        assert ranks[i] == want, (i, o, ranks[i], want)
        if last[i]:
            assert totals[i] == run, (i, totals[i], run)
    # aggregate = counters since the last cell start inside the tile, head = a cell starts inside the tile
    run = [0] * 8
    for i in range(n):
        if first[i]:
            run = [0] * 8
        if octs[i] < 8:
            run[octs[i]] += 1
    assert agg == run and head == any(first)


def test_worst_case_counts_fit_the_fields():
    # 4096 losers of ONE octant in a tile: the 8-bit thread counters reach 16, the 16-bit fields 4096 (bit 12),
    # the flag in bit 31 of the fourth register survives the additions
    for o in (0, 3, 4, 7):
        octs, first, last = [o] * TILE, [0] * TILE, [0] * TILE
        first[5] = 1
        ranks, _, agg, head = kernel_model(octs, first, last, [7] * 8)
        assert ranks[:5] == [7, 8, 9, 10, 11] and ranks[5:] == list(range(TILE - 5))
        assert agg[o] == TILE - 5 and sum(agg) == TILE - 5 and head
