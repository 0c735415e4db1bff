"""Shared test helpers: ctypes binding of the CPU oracle, node-file parser and
canonical comparator.  Test infrastructure only."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
from pathlib import Path
from typing import Dict, Optional

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
ORACLE_SO = ORACLE_DIR / "liboracle.so"

POINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("rgba", "u1", (4,))])
assert POINT_DTYPE.itemsize == 16


def build_oracle() -> Path:
    subprocess.check_call(["make", "-s", "-C", str(ORACLE_DIR)])
    return ORACLE_SO


_orc = None


def oracle_lib():
    global _orc
    if _orc is None:
        build_oracle()
        L = C.CDLL(str(ORACLE_SO))
        L.orc_open.restype = C.c_void_p
        L.orc_open.argtypes = [C.c_char_p, C.c_uint32, C.c_int]
        L.orc_set_config.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_float]
        L.orc_add_file.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.orc_finish.argtypes = [C.c_void_p]
        L.orc_close.argtypes = [C.c_void_p]
        L.orc_number_of_points.restype = C.c_uint64
        L.orc_number_of_points.argtypes = [C.c_void_p]
        L.orc_hierarchies.restype = C.c_uint32
        L.orc_hierarchies.argtypes = [C.c_void_p]
        L.orc_arrivals.restype = C.c_uint64
        L.orc_arrivals.argtypes = [C.c_void_p]
        L.orc_points_by_delta.restype = C.c_int64
        L.orc_points_by_delta.argtypes = [C.c_void_p]
        L.orc_cell_count.restype = C.c_uint64
        L.orc_cell_count.argtypes = [C.c_void_p]
        L.orc_bounding_box.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_dump_nodes.restype = C.c_uint64
        L.orc_dump_nodes.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.orc_metadata_json.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64]
        L.orc_cell_size.restype = C.c_float
        L.orc_cell_size.argtypes = [C.c_uint32, C.c_uint32, C.c_float, C.c_uint32]
        L.orc_cell_index.argtypes = [C.c_float, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_cell_pos.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
        L.orc_hex_from_world.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
        L.orc_hex_to_world.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
        L.orc_hex_d2.restype = C.c_float
        L.orc_hex_d2.argtypes = [C.c_void_p, C.c_float]
        L.orc_format_f32.argtypes = [C.c_float, C.c_char_p, C.c_uint64]
        L.orc_set_lru.argtypes = [C.c_void_p, C.c_uint32]
        L.orc_lru_stats.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _orc = L
    return _orc


def as_points(arr) -> np.ndarray:
    a = np.ascontiguousarray(arr)
    if a.dtype != POINT_DTYPE:
        a = a.view(np.uint8).reshape(-1, 16).view(POINT_DTYPE).reshape(-1)
    return a


def make_points(xyz: np.ndarray, rgba: Optional[np.ndarray] = None) -> np.ndarray:
    xyz = np.asarray(xyz, dtype=np.float32).reshape(-1, 3)
    out = np.zeros(len(xyz), dtype=POINT_DTYPE)
    out["x"], out["y"], out["z"] = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    if rgba is None:
        # colour encodes the input ordinal so duplicates stay distinguishable
        idx = np.arange(len(xyz), dtype=np.uint32)
        out["rgba"] = idx.view(np.uint8).reshape(-1, 4)
    else:
        out["rgba"] = rgba
    return out


class OracleRun:
    """Result of one oracle conversion: {path: node bytes} + metadata."""

    def __init__(self, nodes: Dict[str, bytes], meta_json: str, number_of_points, hierarchies, bbox, arrivals):
        self.nodes = nodes
        self.meta_json = meta_json
        self.number_of_points = number_of_points
        self.hierarchies = hierarchies
        self.bbox = bbox
        self.arrivals = arrivals


def oracle_convert(files, out_dir: Optional[str] = None, config=None, batch_size=10_000, threads=1,
                   dump=True, lru=0) -> OracleRun:
    """files: list of point arrays, each one input file (lib.rs:17-52).  lru > 0: emulate the reference CLI's
    cell residency (converter.rs:92: LRU of `lru` cells spilling whole node files to out_dir); the dump then only
    holds the cells still resident - read the tree from out_dir instead."""
    L = oracle_lib()
    h = L.orc_open(out_dir.encode() if out_dir else None, batch_size, threads)
    assert h, "oracle failed to open (bad metadata.json?)"
    try:
        if config is not None:
            L.orc_set_config(h, int(config[0]), int(config[1]), float(config[2]))
        if lru:
            assert L.orc_set_lru(h, int(lru)) == 0, "LRU emulation needs an output directory and one thread"
        for f in files:
            a = as_points(f)
            rc = L.orc_add_file(h, a.ctypes.data if len(a) else None, len(a))
            assert rc == 0, "oracle: depth >= 32"
        assert L.orc_finish(h) == 0
        nodes = {}
        if dump:
            size = L.orc_dump_nodes(h, None, 0)
            buf = np.empty(size, dtype=np.uint8)
            L.orc_dump_nodes(h, buf.ctypes.data, size)
            nodes = parse_dump(buf)
        mj = C.create_string_buffer(4096)
        L.orc_metadata_json(h, mj, 4096)
        bb = np.zeros(6, dtype=np.float32)
        L.orc_bounding_box(h, bb.ctypes.data)
        run = OracleRun(nodes, mj.value.decode(), L.orc_number_of_points(h), L.orc_hierarchies(h), bb,
                        L.orc_arrivals(h))
        run.points_by_delta = L.orc_points_by_delta(h)   # the GUI caller's count (src/plugins/converter.rs:565-573)
        return run
    finally:
        L.orc_close(h)


def parse_dump(buf: np.ndarray) -> Dict[str, bytes]:
    out = {}
    off = 0
    mv = memoryview(buf)
    n = len(buf)
    while off < n:
        pl = int(np.frombuffer(mv[off:off + 4], dtype="<u4")[0])
        path = bytes(mv[off + 4:off + 4 + pl]).decode()
        il = int(np.frombuffer(mv[off + 4 + pl:off + 12 + pl], dtype="<u8")[0])
        out[path] = bytes(mv[off + 12 + pl:off + 12 + pl + il])
        off += 12 + pl + il
    return out


def read_tree(out_dir) -> Dict[str, bytes]:
    """Read every h_*/c_*.bin under out_dir into {relative path: bytes}."""
    out = {}
    base = Path(out_dir)
    for hdir in sorted(base.glob("h_*")):
        for f in hdir.iterdir():
            if f.suffix == ".bin":
                out[f"{hdir.name}/{f.name}"] = f.read_bytes()
    return out


HEADER_DTYPE = np.dtype([("h", "<u4"), ("idx", "<i4", (3,)), ("total", "<u4"), ("npoints", "<u4"),
                         ("noverflow", "<u4"), ("size", "<f4"), ("sub", "<f4"), ("pos", "<f4", (3,))])
assert HEADER_DTYPE.itemsize == 48


class Node:
    """Parsed node file (layout: point-converter/src/cell.rs:155-229, 280-335)."""

    def __init__(self, data: bytes):
        self.raw = data
        self.header = np.frombuffer(data, dtype=HEADER_DTYPE, count=1)[0]
        g = int(self.header["npoints"])
        off = 48
        self.grid = np.frombuffer(data, dtype=np.uint8, count=16 * g, offset=off).reshape(g, 16)
        off += 16 * g
        ne = data[off]
        off += 1
        self.entries = {}  # child index tuple -> None (flushed) | (n,16) uint8 array in list order
        for _ in range(ne):
            key = tuple(int(v) for v in np.frombuffer(data, dtype="<i4", count=3, offset=off))
            cnt = int(np.frombuffer(data, dtype="<u4", count=1, offset=off + 12)[0])
            off += 16
            if cnt == 0:
                self.entries[key] = None
            else:
                self.entries[key] = np.frombuffer(data, dtype=np.uint8, count=16 * cnt, offset=off).reshape(cnt, 16)
                off += 16 * cnt
        assert off == len(data), f"trailing bytes in node: {len(data) - off}"

    def canonical(self):
        """Order-free form: header bytes, grid as a sorted multiset of 16-byte records,
        overflow as {child: None | list bytes in order}."""
        g = self.grid
        if len(g):
            order = np.lexsort(g.T[::-1])
            g = g[order]
        return (self.raw[:48], g.tobytes(),
                {k: (None if v is None else v.tobytes()) for k, v in self.entries.items()})


def compare_trees(a: Dict[str, bytes], b: Dict[str, bytes], label_a="oracle", label_b="gpu", max_report=5):
    """Assert two conversions are identical up to the in-file orders that the
    reference leaves to HashMap iteration."""
    ka, kb = set(a), set(b)
    only_a, only_b = sorted(ka - kb), sorted(kb - ka)
    assert not only_a and not only_b, (
        f"node sets differ: only in {label_a}: {only_a[:max_report]} ({len(only_a)}); "
        f"only in {label_b}: {only_b[:max_report]} ({len(only_b)})")
    bad = []
    for k in sorted(ka):
        if a[k] == b[k]:
            continue
        ca, cb = Node(a[k]).canonical(), Node(b[k]).canonical()
        if ca != cb:
            what = []
            if ca[0] != cb[0]:
                what.append(f"header {Node(a[k]).header} vs {Node(b[k]).header}")
            if ca[1] != cb[1]:
                what.append("grid set")
            if ca[2] != cb[2]:
                what.append(f"overflow keys {sorted(ca[2])} vs {sorted(cb[2])}"
                            if set(ca[2]) != set(cb[2]) else "overflow lists")
            bad.append((k, "; ".join(what)))
    assert not bad, f"{len(bad)} nodes differ, first: {bad[:max_report]}"


def tree_points(nodes: Dict[str, bytes]) -> np.ndarray:
    """All stored points (grid then overflow) as an (n,16) uint8 array."""
    parts = []
    for data in nodes.values():
        nd = Node(data)
        parts.append(nd.grid)
        for v in nd.entries.values():
            if v is not None:
                parts.append(v)
    if not parts:
        return np.zeros((0, 16), dtype=np.uint8)
    return np.concatenate(parts, axis=0)


def sort_records(rec: np.ndarray) -> np.ndarray:
    rec = np.ascontiguousarray(rec).view(np.uint8).reshape(-1, 16)
    if len(rec) == 0:
        return rec
    return rec[np.lexsort(rec.T[::-1])]


def records_digest(rec: np.ndarray):
    """Order-independent digest of 16-byte records: (count, sum and xor of a 64-bit mix of both halves)."""
    w = np.ascontiguousarray(rec).view(np.uint64).reshape(-1, 2)
    with np.errstate(over="ignore"):
        h = (w[:, 0] * np.uint64(0x9E3779B97F4A7C15)) ^ (w[:, 1] * np.uint64(0xC2B2AE3D27D4EB4F))
        h ^= h >> np.uint64(29)
        h *= np.uint64(0xBF58476D1CE4E5B9)
        h ^= h >> np.uint64(32)
        return len(w), int(h.sum(dtype=np.uint64)), int(np.bitwise_xor.reduce(h)) if len(h) else 0


def combine_digests(a, b):
    return a[0] + b[0], (a[1] + b[1]) & 0xFFFFFFFFFFFFFFFF, a[2] ^ b[2]


def check_tree_cells(nodes: Dict[str, bytes], limit=5000, max_cell=1000.0):
    """Size-independent properties of a converted tree: header accounting, overflow lists within the limit, every
    stored point inside its node's cell under the reference's f32 division (metadata.rs:100-102).  Returns the digest
    of all stored records."""
    import re
    total = (0, 0, 0)
    for path, data in nodes.items():
        m = re.fullmatch(r"h_(\d+)/c_(-?\d+)_(-?\d+)_(-?\d+)\.bin", path)
        h, idx = int(m.group(1)), np.array([int(m.group(k)) for k in (2, 3, 4)], dtype=np.int64)
        nd = Node(data)
        lists = [v for v in nd.entries.values() if v is not None]
        assert all(len(v) <= limit for v in lists), f"{path}: overflow list longer than the limit"
        stored = np.concatenate([nd.grid] + lists, axis=0) if lists else nd.grid
        assert int(nd.header["total"]) == len(stored) and int(nd.header["npoints"]) == len(nd.grid), path
        size = np.float32(max_cell) / np.float32(2 ** h)
        xyz = np.ascontiguousarray(stored[:, :12]).view(np.float32).reshape(-1, 3)
        cell = np.floor(xyz / size).astype(np.int64)
        assert np.all(cell == idx[None, :]), f"{path}: a stored point lies outside the node's cell"
        total = combine_digests(total, records_digest(stored))
    return total


def oracle_cull(params, hier, loaded, far, depth, pts) -> np.ndarray:
    """CPU restatement of compute.wgsl:81-205 (oracle/cull_oracle.cpp).  params: pcb.PcbCullParams (same layout);
    hier (nh, 2) f32; loaded (m, 4) i32 sorted; far (k,) f32; depth (h, w) f32; pts: point records.  Returns the kept
    records (input order)."""
    L = oracle_lib()
    L.orc_cull_lod.restype = C.c_uint64
    L.orc_cull_lod.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                               C.c_uint64, C.c_void_p]
    hier = np.ascontiguousarray(hier, dtype=np.float32)
    loaded = np.ascontiguousarray(loaded, dtype=np.int32).reshape(-1, 4)
    far = np.ascontiguousarray(far, dtype=np.float32)
    depth = np.ascontiguousarray(depth, dtype=np.float32)
    a = as_points(pts)
    out = np.zeros(max(len(a), 1), dtype=POINT_DTYPE)
    k = L.orc_cull_lod(C.addressof(params), hier.ctypes.data, loaded.ctypes.data, len(loaded), far.ctypes.data,
                       depth.ctypes.data, a.ctypes.data, len(a), out.ctypes.data)
    return out[:k]


def load_pcb():
    """Import point-cloud_b200/pcb.py (the directory name is not a Python identifier)."""
    import importlib.util
    name = "point_cloud_b200_pcb"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, ROOT / "point-cloud_b200" / "pcb.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod
