"""tests/bulk_model.py — TEST INFRASTRUCTURE ONLY.

numpy model of the BULK, level-synchronous formulation that the CUDA path
(point-cloud_b200/csrc) implements.  It exists so that the algorithm itself
(arrival keys, emission order, the per-(cell, child) flush rule, merge-mode
initial state — SURVEY.md §8 N3) can be validated against the sequential oracle
on the CPU, kernel step by kernel step, before any GPU time is spent.  The CUDA
host driver follows this file stage by stage (names match build.cu).

Reference semantics being reproduced (paths relative to /root/reference):
  point-converter/src/converter.rs:32-139, cell.rs:70-153, hex.rs:17-85,
  metadata.rs:90-106, lib.rs:31-52.
"""
from __future__ import annotations

import struct
from typing import Dict, List, Optional

import numpy as np

f32 = np.float32
SQRT_3 = f32(1.73205080757)  # hex.rs:3
NONE = -1


def sat_i32(v: np.ndarray) -> np.ndarray:
    """Rust `f32 as i32` (saturating, NaN->0), vectorised."""
    v = np.asarray(v, dtype=np.float32)
    out = np.zeros(v.shape, dtype=np.int64)
    ok = ~np.isnan(v)
    hi = ok & (v >= f32(2147483648.0))
    lo = ok & (v <= f32(-2147483648.0))
    mid = ok & ~hi & ~lo
    out[hi] = 2**31 - 1
    out[lo] = -(2**31)
    out[mid] = np.trunc(v[mid]).astype(np.int64)
    return out


def cell_size(max_cell_size, h):  # metadata.rs:91-93
    return f32(max_cell_size) / f32(2**h)


def cell_index(xyz: np.ndarray, cs) -> np.ndarray:  # metadata.rs:100-102
    return sat_i32(np.floor(xyz / f32(cs)))


def hex_offset_key(xyz: np.ndarray, R) -> np.ndarray:
    """hex.rs:67-85 then :45-51, vectorised; returns (n,3) int64 (X, Y, Z)."""
    R = f32(R)
    x = xyz[:, 0] / (R * SQRT_3)
    y = xyz[:, 1] / ((-R) * SQRT_3)
    t = SQRT_3 * y + f32(1.0)
    t1 = np.floor(t + x)
    t2 = t - x
    t3 = f32(2.0) * x + f32(1.0)
    qf = (t1 + t3) / f32(3.0)
    rf = (t1 + t2) / f32(3.0)
    q = sat_i32(np.floor(qf))
    r = -sat_i32(np.floor(rf))
    r = ((r + 2**31) % 2**32) - 2**31  # wrapping negate of i32::MIN
    hz = sat_i32(xyz[:, 2] / R)
    e = r - (r & 1)
    X = q + np.where(e < 0, -((-e) // 2), e // 2)  # Rust `/` truncates; e is even so exact
    X = ((X + 2**31) % 2**32) - 2**31
    return np.stack([X, r, hz], axis=1)


def hex_centre(key: np.ndarray, R) -> np.ndarray:
    """hex.rs:18-24 then :55-65."""
    R = f32(R)
    X, Y, Z = key[:, 0], key[:, 1], key[:, 2]
    e = Y - (Y & 1)
    q = X - np.where(e < 0, -((-e) // 2), e // 2)
    q = ((q + 2**31) % 2**32) - 2**31
    qf, rf, hf = q.astype(np.float32), Y.astype(np.float32), Z.astype(np.float32)
    cx = R * (SQRT_3 * qf + (SQRT_3 / f32(2.0)) * rf)
    cy = ((R * f32(3.0)) / f32(2.0)) * rf
    cz = hf * R
    return np.stack([cx, cy, cz], axis=1).astype(np.float32)


def dist2(c: np.ndarray, p: np.ndarray) -> np.ndarray:  # glam distance_squared
    with np.errstate(over="ignore", invalid="ignore"):
        d = (c - p).astype(np.float32)
        return (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]


class ExistingNode:
    """Initial state of one cell read back from its node file (merge mode)."""

    def __init__(self, grid: np.ndarray, entries: Dict[tuple, Optional[np.ndarray]]):
        self.grid = grid          # (g,16) uint8
        self.entries = entries    # child index -> None | (n,16) uint8 list in order


def _parse_node(data: bytes) -> ExistingNode:
    """cell.rs:183-229 — node file -> initial state."""
    g = struct.unpack_from("<I", data, 20)[0]
    grid = np.frombuffer(data, dtype=np.uint8, count=16 * g, offset=48).reshape(g, 16).copy()
    off = 48 + 16 * g
    ne = data[off]
    off += 1
    entries = {}
    for _ in range(ne):
        key = struct.unpack_from("<3i", data, off)
        cnt = struct.unpack_from("<I", data, off + 12)[0]
        off += 16
        if cnt == 0:
            entries[key] = None
        else:
            entries[key] = np.frombuffer(data, dtype=np.uint8, count=16 * cnt, offset=off).reshape(cnt, 16).copy()
            off += 16 * cnt
    return ExistingNode(grid, entries)


def _xyz(rec: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(rec[:, :12]).view(np.float32).reshape(-1, 3)


def _seg_prefix_min_excl(bits: np.ndarray, gid: np.ndarray, head: np.ndarray) -> np.ndarray:
    """Exclusive prefix-min of `bits` (uint32 as int64) inside groups `gid` (non-decreasing)."""
    v = bits - gid * (1 << 32)
    cm = np.minimum.accumulate(v) + gid * (1 << 32)
    ex = np.empty_like(cm)
    ex[0] = 1 << 40
    ex[1:] = cm[:-1]
    ex[head] = 1 << 40
    return ex


def flush_rule(b_e: np.ndarray, m0: int, state: str, L: int):
    """cell.rs:115-149 for ONE (cell, child) bucket, closed form.

    b_e: batch ordinals of the NEW losers in emission order (non-decreasing).
    state: 'absent' | 'some' (m0 existing list entries) | 'none'.
    Returns (flushed: bool, b_flush: int)  — b_flush is the batch whose call delivers the list
    (meaningless when state == 'none', where every call delivers its own points)."""
    n = len(b_e)
    if state == "none":
        return True, -1
    if state == "some":
        if m0 + n < L:
            return False, 0
        k = max(L - 1 - m0, 0)  # first new element whose arrival makes the list reach L
        return True, int(b_e[k])
    # absent: first call keeps when n_1 <= L (cell.rs:118); later calls flush when len >= L (:137)
    if n < L:
        return False, 0
    bstar = int(b_e[L - 1])
    if bstar == int(b_e[0]) and (n == L or int(b_e[L]) != bstar):
        # the first call delivered exactly L losers: kept as Some(L); the next call (if any) flushes
        if n == L:
            return False, 0
        return True, int(b_e[L])
    return True, bstar


def bulk_build(files: List[np.ndarray], config=(5000, 96, 1000.0), batch_size=10_000,
               existing: Optional[Dict[str, bytes]] = None, meta0: Optional[dict] = None):
    """Level-synchronous build.  files: list of (n,16)-byte point arrays (one per input file).
    existing/meta0: merge mode (node files + metadata of the previous run).
    Returns (nodes {path: bytes} of every node touched by this run, meta dict)."""
    L, dim, max_cs = int(config[0]), int(config[1]), f32(config[2])
    existing = existing or {}
    meta = dict(number_of_points=0, hierarchies=0, bbox=None)
    if meta0:
        meta.update(meta0)

    # ---- stage 0: level-0 arrivals: (cell0, input index) order, b = global batch ordinal (lib.rs:31-52)
    recs, bs = [], []
    b0 = 0
    for f in files:
        r = np.ascontiguousarray(f).view(np.uint8).reshape(-1, 16)
        n = len(r)
        recs.append(r)
        bs.append(b0 + np.arange(n, dtype=np.int64) // batch_size)
        b0 += max(1, -(-n // batch_size))  # an empty file still yields one (empty) batch
    rec = np.concatenate(recs) if recs else np.zeros((0, 16), np.uint8)
    b = np.concatenate(bs) if bs else np.zeros(0, np.int64)
    xyz = _xyz(rec)
    if len(rec):
        lo, hi = xyz.min(axis=0), xyz.max(axis=0)
        if meta["number_of_points"] == 0 or meta["bbox"] is None:
            meta["bbox"] = (lo, hi)
        else:
            meta["bbox"] = (np.minimum(meta["bbox"][0], lo), np.maximum(meta["bbox"][1], hi))
    meta["number_of_points"] += len(rec)
    nodes: Dict[str, bytes] = {}
    if not files:
        return nodes, meta
    meta["hierarchies"] = max(meta["hierarchies"], 1)  # level 0 is visited by every batch, even empty

    ci = cell_index(xyz, cell_size(max_cs, 0))
    cells, cell_of = np.unique(ci, axis=0, return_inverse=True)
    cell_of = cell_of.reshape(-1)
    order = np.argsort(cell_of, kind="stable")
    rec, b, cell_of = rec[order], b[order], cell_of[order]

    h = 0
    arrivals_total = 0
    while len(rec):
        assert h < 32, "octree deeper than 31 levels"
        meta["hierarchies"] = max(meta["hierarchies"], h + 1)
        arrivals_total += len(rec)
        S = cell_size(max_cs, h)
        sub = S / f32(dim)
        R = sub / f32(2.0)
        ncell = len(cells)

        # ---- merge mode: incumbents (existing grid winners) go in front of each cell's arrivals
        ex_nodes: Dict[int, ExistingNode] = {}
        for c in range(ncell):
            p = "h_%d/c_%d_%d_%d.bin" % (h, cells[c][0], cells[c][1], cells[c][2])
            if p in existing:
                ex_nodes[c] = _parse_node(existing[p])
        if ex_nodes:
            parts_r, parts_b, parts_c, parts_i = [rec], [b], [cell_of], [np.zeros(len(rec), bool)]
            for c, en in ex_nodes.items():
                parts_r.append(en.grid)
                parts_b.append(np.full(len(en.grid), -1, np.int64))
                parts_c.append(np.full(len(en.grid), c, np.int64))
                parts_i.append(np.ones(len(en.grid), bool))
            rec = np.concatenate(parts_r)
            b = np.concatenate(parts_b)
            cell_of = np.concatenate(parts_c)
            is_inc = np.concatenate(parts_i)
            o = np.lexsort((b, cell_of))  # stable: incumbents (b=-1) first inside each cell
            rec, b, cell_of, is_inc = rec[o], b[o], cell_of[o], is_inc[o]
        else:
            is_inc = np.zeros(len(rec), bool)
        xyz = _xyz(rec)
        n = len(rec)
        pos = np.arange(n, dtype=np.int64)  # arrival position == proxy of tau inside a cell

        # ---- K_hexkey: hex key + d2 per arrival
        key = hex_offset_key(xyz, R)
        d2 = dist2(hex_centre(key, R), xyz)
        bits = d2.view(np.uint32).astype(np.int64)

        # ---- radix sort by (cell, Z, Y, X), stable in arrival position
        srt = np.lexsort((pos, key[:, 0], key[:, 1], key[:, 2], cell_of))
        skey = np.concatenate([cell_of[srt, None], key[srt]], axis=1)
        head = np.ones(n, bool)
        head[1:] = np.any(skey[1:] != skey[:-1], axis=1)
        gid = np.cumsum(head) - 1

        # ---- K_contest: strict prefix minima = records; last record wins; losers get trigger
        sb = bits[srt]
        ex = _seg_prefix_min_excl(sb, gid, head)
        is_rec = head | (sb < ex)
        idx = np.arange(n, dtype=np.int64)
        nxt_src = np.where(is_rec, idx, n + 1)[::-1]
        nxt_incl = np.minimum.accumulate(nxt_src)[::-1]          # next record at or after i
        nxt = np.full(n, n + 1, np.int64)
        nxt[:-1] = nxt_incl[1:]                                   # next record strictly after i
        tail_gid = np.append(gid, -1)
        nxt_same = (nxt <= n - 1) & (tail_gid[np.minimum(nxt, n)] == gid)
        winner = is_rec & ~nxt_same
        trig_sorted = np.where(is_rec, nxt, idx)                  # sorted index of the trigger
        loser = ~winner
        emitted = np.full(n, NONE, np.int64)                      # arrival pos of trigger -> loser pos
        emitted[srt[trig_sorted[loser]]] = srt[loser]

        # ---- emission order = trigger position order (cell-major, then tau)
        trig_pos = np.nonzero(emitted != NONE)[0]
        lsrc = emitted[trig_pos]
        l_b = b[trig_pos]
        l_cell = cell_of[trig_pos]
        assert np.all(cell_of[lsrc] == l_cell)
        child = cell_index(xyz[lsrc], cell_size(max_cs, h + 1))
        d = child - 2 * cells[l_cell]
        assert np.all((d >= 0) & (d <= 1)), "child index does not nest (N1)"
        oct_ = d[:, 0] * 4 + d[:, 1] * 2 + d[:, 2]
        bucket = l_cell * 8 + oct_
        bo = np.argsort(bucket, kind="stable")                    # segmented 8-way stable split
        lsrc, l_b, bucket = lsrc[bo], l_b[bo], bucket[bo]

        # ---- K_bucket: flush decision per (cell, child)
        ub, ustart, ucount = np.unique(bucket, return_index=True, return_counts=True)
        next_parts = []  # (child cell index, rec, b)
        new_entries: Dict[int, Dict[tuple, Optional[np.ndarray]]] = {}
        for k in range(len(ub)):
            c, o8 = int(ub[k]) // 8, int(ub[k]) % 8
            cidx = tuple(int(v) for v in (2 * cells[c] + np.array([(o8 >> 2) & 1, (o8 >> 1) & 1, o8 & 1])))
            sl = slice(ustart[k], ustart[k] + ucount[k])
            new_rec, new_b = rec[lsrc[sl]], l_b[sl]
            old = ex_nodes[c].entries.get(cidx, "absent") if c in ex_nodes else "absent"
            if isinstance(old, str):
                state, m0, old_rec = "absent", 0, np.zeros((0, 16), np.uint8)
            elif old is None:
                state, m0, old_rec = "none", 0, np.zeros((0, 16), np.uint8)
            else:
                state, m0, old_rec = "some", len(old), old
            if L == 0 and state == "absent":
                flushed, bf = True, int(new_b[0])
            else:
                flushed, bf = flush_rule(new_b, m0, state, L)
            ent = new_entries.setdefault(c, {})
            if not flushed:
                ent[cidx] = np.concatenate([old_rec, new_rec])
            else:
                ent[cidx] = None
                drec = np.concatenate([old_rec, new_rec])
                db = np.concatenate([np.full(m0, -1, np.int64), new_b])
                next_parts.append((cidx, drec, np.maximum(db, bf)))

        # ---- K_assemble: node images of every cell that received arrivals
        w_sorted = np.nonzero(winner)[0]
        w_cell = skey[w_sorted, 0]
        w_start = np.searchsorted(w_cell, np.arange(ncell), side="left")
        w_end = np.searchsorted(w_cell, np.arange(ncell), side="right")
        for c in range(ncell):
            entries = dict(ex_nodes[c].entries) if c in ex_nodes else {}
            entries.update(new_entries.get(c, {}))
            grid = rec[srt[w_sorted[w_start[c]:w_end[c]]]]
            nov = sum(len(v) for v in entries.values() if v is not None)
            cs = cells[c]
            half = S / f32(2.0)
            cpos = [f32(int(cs[a])) * S + half for a in range(3)]
            out = [struct.pack("<I3i3I", h, int(cs[0]), int(cs[1]), int(cs[2]), len(grid) + nov, len(grid), nov),
                   struct.pack("<5f", S, sub, *cpos), grid.tobytes(), struct.pack("<B", len(entries))]
            for cidx in sorted(entries):
                v = entries[cidx]
                out.append(struct.pack("<3i", *cidx))
                out.append(struct.pack("<I", 0 if v is None else len(v)))
                if v is not None:
                    out.append(v.tobytes())
            nodes["h_%d/c_%d_%d_%d.bin" % (h, cs[0], cs[1], cs[2])] = b"".join(out)

        # ---- next level: delivered buckets, already in (child cell, tau') order
        if not next_parts:
            break
        next_parts.sort(key=lambda t: t[0])
        cells = np.array([t[0] for t in next_parts], dtype=np.int64)
        rec = np.concatenate([t[1] for t in next_parts])
        b = np.concatenate([t[2] for t in next_parts])
        cell_of = np.concatenate([np.full(len(t[1]), i, np.int64) for i, t in enumerate(next_parts)])
        h += 1
    meta["arrivals"] = arrivals_total
    return nodes, meta
