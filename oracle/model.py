"""oracle/model.py — TEST INFRASTRUCTURE ONLY.

Second, independent restatement of the reference's octree/LOD construction in
pure Python + numpy.float32 scalars, written directly from the Rust sources
(not from pc_oracle.cpp).  Its only job is to pin the C++ oracle: both must
produce bit-identical node contents on small inputs (tests/test_oracle_model.py).

PARITY UNPINNED by the reference's own tests (it has none for this path,
SURVEY.md §4); the reference cannot be run here (no Rust toolchain).

Every function cites the reference file:line (relative to /root/reference).
All arithmetic is np.float32 op by op (numpy rounds each scalar op to binary32,
no FMA), evaluated left to right as in the Rust source.
"""
from __future__ import annotations

import struct
from typing import Dict, List, Optional, Tuple

import numpy as np

f32 = np.float32
SQRT_3 = f32(1.73205080757)  # point-converter/src/hex.rs:3

Point = Tuple[np.float32, np.float32, np.float32, bytes]  # (x, y, z, rgba)


def sat_i32(v: np.float32) -> int:
    """Rust `f32 as i32`: saturating, NaN -> 0."""
    if np.isnan(v):
        return 0
    if v >= f32(2147483648.0):
        return 2**31 - 1
    if v <= f32(-2147483648.0):
        return -(2**31)
    return int(v)  # truncation toward zero


class Config:
    """point-converter/src/metadata.rs:67-113"""

    def __init__(self, limit=5000, dim=96, max_cell_size=1000.0):
        self.limit = int(limit)
        self.dim = int(dim)
        self.max_cell_size = f32(max_cell_size)

    def cell_size(self, h: int) -> np.float32:  # metadata.rs:91-93
        return self.max_cell_size / f32(2**h)

    def sub_cell_size(self, cell_size) -> np.float32:  # metadata.rs:95-97
        return cell_size / f32(self.dim)

    def cell_index(self, p, cell_size):  # metadata.rs:100-102
        return tuple(sat_i32(np.floor(p[a] / cell_size)) for a in range(3))

    def cell_pos(self, idx, cell_size):  # metadata.rs:104-106
        half = cell_size / f32(2.0)
        return tuple(f32(idx[a]) * cell_size + half for a in range(3))


def to_axial(o):  # hex.rs:18-24
    x, y, z = o
    t = y - (y & 1)
    return (x - int(t / 2), y, z)  # t is even: exact


def to_offset(a):  # hex.rs:45-51
    q, r, h = a
    t = r - (r & 1)
    return (q + int(t / 2), r, h)


def axial_to_world(a, R):  # hex.rs:55-65
    q, r, h = f32(a[0]), f32(a[1]), f32(a[2])
    x = R * (SQRT_3 * q + SQRT_3 / f32(2.0) * r)
    y = R * f32(3.0) / f32(2.0) * r
    z = h * R
    return (x, y, z)


def axial_from_world(p, R):  # hex.rs:67-85
    x = p[0] / (R * SQRT_3)
    y = p[1] / (-R * SQRT_3)
    t = SQRT_3 * y + f32(1.0)
    temp1 = np.floor(t + x)
    temp2 = t - x
    temp3 = f32(2.0) * x + f32(1.0)
    qf = (temp1 + temp3) / f32(3.0)
    rf = (temp1 + temp2) / f32(3.0)
    q = sat_i32(np.floor(qf))
    r = -sat_i32(np.floor(rf))
    h = sat_i32(p[2] / R)
    return (q, r, h)


def offset_from_world(p, R):  # hex.rs:32-34
    return to_offset(axial_from_world(p, R))


def offset_to_world(o, R):  # hex.rs:28-30
    return axial_to_world(to_axial(o), R)


def distance_squared(a, b):  # glam 0.27.0 Vec3::distance_squared
    dx, dy, dz = a[0] - b[0], a[1] - b[1], a[2] - b[2]
    return dx * dx + dy * dy + dz * dz


class Cell:
    """point-converter/src/cell.rs:33-229"""

    def __init__(self, h, idx, sub_cell_size, size, pos):
        self.h, self.idx = h, idx
        self.total = 0
        self.npoints = 0
        self.noverflow = 0
        self.size, self.sub_cell_size, self.pos = size, sub_cell_size, pos
        self.grid: Dict[Tuple[int, int, int], Point] = {}
        self.overflow: Dict[Tuple[int, int, int], Optional[List[Point]]] = {}

    def add_point(self, p: Point) -> Optional[Point]:  # cell.rs:70-94
        R = self.sub_cell_size / f32(2.0)
        key = offset_from_world(p, R)
        if key in self.grid:
            centre = offset_to_world(key, R)
            old = distance_squared(centre, self.grid[key])
            new = distance_squared(centre, p)
            if new < old:
                out = self.grid[key]
                self.grid[key] = p
                return out
            return p
        self.grid[key] = p
        self.total += 1
        self.npoints += 1
        return None

    def add_points(self, pts: List[Point]) -> List[Point]:  # cell.rs:96-106
        out = []
        for p in pts:
            l = self.add_point(p)
            if l is not None:
                out.append(l)
        return out

    def add_points_in_overflow(self, groups, cfg: Config):  # cell.rs:108-153
        remaining = {}
        for cidx, pts in groups.items():
            if cidx not in self.overflow:
                if len(pts) <= cfg.limit:
                    self.total += len(pts)
                    self.noverflow += len(pts)
                    self.overflow[cidx] = list(pts)
                else:
                    remaining[cidx] = list(pts)
                    self.overflow[cidx] = None
            elif self.overflow[cidx] is None:
                remaining[cidx] = list(pts)
            else:
                lst = self.overflow[cidx]
                old_len = len(lst)
                lst.extend(pts)
                if len(lst) < cfg.limit:
                    self.total += len(pts)
                    self.noverflow += len(pts)
                else:
                    self.total -= old_len
                    self.noverflow -= old_len
                    remaining[cidx] = lst
                    self.overflow[cidx] = None
        return remaining

    def to_bytes(self) -> bytes:  # cell.rs:155-181, 280-298 (canonical order, see pc_oracle.cpp)
        out = [struct.pack("<I3i3I", self.h, *self.idx, self.total, self.npoints, self.noverflow)]
        out.append(struct.pack("<5f", self.size, self.sub_cell_size, *self.pos))
        for key in sorted(self.grid, key=lambda k: (k[2], k[1], k[0])):
            p = self.grid[key]
            out.append(struct.pack("<3f", p[0], p[1], p[2]) + p[3])
        out.append(struct.pack("<B", len(self.overflow)))
        for cidx in sorted(self.overflow):
            lst = self.overflow[cidx]
            out.append(struct.pack("<3i", *cidx))
            if lst is None:
                out.append(struct.pack("<I", 0))
            else:
                out.append(struct.pack("<I", len(lst)))
                for p in lst:
                    out.append(struct.pack("<3f", p[0], p[1], p[2]) + p[3])
        return b"".join(out)


def group_points(pts: List[Point], h: int, cfg: Config):  # converter.rs:32-47
    cs = cfg.cell_size(h)
    m: Dict[Tuple[int, int, int], List[Point]] = {}
    for p in pts:
        m.setdefault(cfg.cell_index(p, cs), []).append(p)
    return m


class Converter:
    """point-converter/src/converter.rs:72-246, in RAM, fresh output only."""

    def __init__(self, cfg: Config, batch_size=10_000):
        self.cfg = cfg
        self.batch_size = batch_size  # lib.rs:32
        self.cells: Dict[Tuple[int, Tuple[int, int, int]], Cell] = {}
        self.number_of_points = 0
        self.hierarchies = 0
        self.bbox = None

    def get_cell(self, h, idx) -> Cell:  # converter.rs:160-207
        key = (h, idx)
        if key not in self.cells:
            cs = self.cfg.cell_size(h)
            self.cells[key] = Cell(h, idx, self.cfg.sub_cell_size(cs), cs, self.cfg.cell_pos(idx, cs))
        return self.cells[key]

    def add_points_batch(self, pts: List[Point]):  # converter.rs:96-112
        if pts:
            xs = np.array([[p[0], p[1], p[2]] for p in pts], dtype=np.float32)
            lo, hi = xs.min(axis=0), xs.max(axis=0)
            if self.number_of_points == 0:
                self.bbox = (lo, hi)
            else:
                self.bbox = (np.minimum(self.bbox[0], lo), np.maximum(self.bbox[1], hi))
        self.number_of_points += len(pts)
        grouped = group_points(pts, 0, self.cfg)
        h = 0
        while True:  # converter.rs:114-139
            if self.hierarchies <= h:
                self.hierarchies += 1
            nxt: Dict[Tuple[int, int, int], List[Point]] = {}
            for idx, cpts in grouped.items():
                cell = self.get_cell(h, idx)
                losers = cell.add_points(cpts)  # converter.rs:62-70
                remaining = cell.add_points_in_overflow(group_points(losers, h + 1, self.cfg), self.cfg)
                for k, v in remaining.items():  # merge_point_maps, converter.rs:49-60
                    nxt.setdefault(k, []).extend(v)
            if not nxt:
                break
            grouped = nxt
            h += 1

    def add_file(self, arr: np.ndarray):
        """arr: structured/byte array of 16-byte records.  lib.rs:31-52."""
        raw = np.ascontiguousarray(arr).view(np.uint8).reshape(-1, 16)
        xyz = raw[:, :12].copy().view(np.float32).reshape(-1, 3)
        pts = [(xyz[i, 0], xyz[i, 1], xyz[i, 2], raw[i, 12:16].tobytes()) for i in range(len(raw))]
        off = 0
        while True:
            batch = pts[off:off + self.batch_size]
            self.add_points_batch(batch)
            off += len(batch)
            if off >= len(pts):
                break

    def nodes(self) -> Dict[str, bytes]:
        return {
            "h_%d/c_%d_%d_%d.bin" % (h, idx[0], idx[1], idx[2]): c.to_bytes()
            for (h, idx), c in self.cells.items()
        }
