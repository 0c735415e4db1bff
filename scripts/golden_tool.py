#!/usr/bin/env python
"""golden_tool.py - the portable half of scripts/make_golden.sh (numpy only, no torch, no CUDA).

    golden_tool.py ply    <name> <out.ply>                      dump the named synthetic cloud as binary-LE PLY
    golden_tool.py digest <name> <in.ply> <converted dir> <out.json>   canonical per-node SHA-256 of a converted tree

The clouds follow SURVEY.md 8d's counter-based generator: u = splitmix64(seed ^ (i * 0x9E3779B97F4A7C15 + stream)),
uniform f32 = (u >> 40) * 2^-24 - integer arithmetic up to the final int -> float conversion, so the same bytes come
out on every machine.  `cloud(name)` is also what tests/test_golden.py feeds the oracle and the CUDA path.
"""
import hashlib
import json
import re
import sys
from pathlib import Path

import numpy as np

POINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("rgba", "u1", (4,))])
GOLDEN = {
    # name: (points, seed, scale xyz, offset xyz)
    "c1_cube_1m": (1_000_000, 1, (1.0, 1.0, 1.0), (0.0, 0.0, 0.0)),                    # BASELINE configs[0]
    "multiroot_200k": (200_000, 7, (3000.0, 2000.0, 50.0), (-1000.0, -1000.0, 0.0)),   # 3 x 2 x 1 level-0 cells
}


def _splitmix64(x):
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def _stream(n, seed, stream):
    with np.errstate(over="ignore"):
        i = np.arange(n, dtype=np.uint64)
        return _splitmix64(np.uint64(seed) ^ (i * np.uint64(0x9E3779B97F4A7C15) + np.uint64(stream)))


def cloud(name):
    n, seed, scale, offset = GOLDEN[name]
    out = np.zeros(n, dtype=POINT_DTYPE)
    for a, f in enumerate("xyz"):
        u = (_stream(n, seed, a + 1) >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)
        out[f] = u * np.float32(scale[a]) + np.float32(offset[a])     # two IEEE binary32 operations
    c = _stream(n, seed, 4)
    out["rgba"][:, 0] = (c & np.uint64(0xFF)).astype(np.uint8)
    out["rgba"][:, 1] = ((c >> np.uint64(8)) & np.uint64(0xFF)).astype(np.uint8)
    out["rgba"][:, 2] = ((c >> np.uint64(16)) & np.uint64(0xFF)).astype(np.uint8)
    out["rgba"][:, 3] = 255   # the PLY carries no alpha: point.rs:105-118 fills in 255
    return out


def write_ply(path, pts):
    hdr = ("ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty float x\nproperty float y\n"
           "property float z\nproperty uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n" % len(pts))
    rec = np.zeros(len(pts), dtype=np.dtype([("xyz", "<f4", (3,)), ("rgb", "u1", (3,))]))
    rec["xyz"][:, 0], rec["xyz"][:, 1], rec["xyz"][:, 2] = pts["x"], pts["y"], pts["z"]
    rec["rgb"] = pts["rgba"][:, :3]
    with open(path, "wb") as f:
        f.write(hdr.encode())
        f.write(rec.tobytes())


def canonical_node_sha(data: bytes) -> str:
    """SHA-256 of the order-free form of one node file (layout: cell.rs:155-229): header, grid records sorted
    bytewise, overflow entries sorted by child index, lists in file order (their order is semantic)."""
    npoints = int(np.frombuffer(data, dtype="<u4", count=1, offset=20)[0])
    grid = np.frombuffer(data, dtype=np.uint8, count=16 * npoints, offset=48).reshape(npoints, 16)
    if npoints:
        grid = grid[np.lexsort(grid.T[::-1])]
    off = 48 + 16 * npoints
    ne = data[off]
    off += 1
    entries = []
    for _ in range(ne):
        key = tuple(int(v) for v in np.frombuffer(data, dtype="<i4", count=3, offset=off))
        cnt = int(np.frombuffer(data, dtype="<u4", count=1, offset=off + 12)[0])
        entries.append((key, bytes(data[off:off + 16 + 16 * cnt])))
        off += 16 + 16 * cnt
    assert off == len(data), "trailing bytes in node file"
    h = hashlib.sha256()
    h.update(bytes(data[:48]))
    h.update(grid.tobytes())
    h.update(bytes([ne]))
    for _, blob in sorted(entries):
        h.update(blob)
    return h.hexdigest()


def tree_digest(nodes: dict) -> dict:
    return {k: canonical_node_sha(v) for k, v in sorted(nodes.items())}


def read_tree(out_dir) -> dict:
    out = {}
    for hdir in sorted(Path(out_dir).glob("h_*")):
        for f in hdir.iterdir():
            if re.fullmatch(r"c_-?\d+_-?\d+_-?\d+\.bin", f.name):
                out[f"{hdir.name}/{f.name}"] = f.read_bytes()
    return out


def main(argv):
    if len(argv) >= 4 and argv[1] == "ply":
        write_ply(argv[3], cloud(argv[2]))
        return 0
    if len(argv) >= 6 and argv[1] == "digest":
        name, ply, out_dir, dst = argv[2:6]
        doc = {"name": name, "ply_sha256": hashlib.sha256(Path(ply).read_bytes()).hexdigest(),
               "producer": "Seiichi-Yahiro/point-cloud point-converter (cargo build --release), default settings",
               "metadata_json": (Path(out_dir) / "metadata.json").read_text(),
               "nodes": tree_digest(read_tree(out_dir))}
        Path(dst).write_text(json.dumps(doc, indent=1, sort_keys=True) + "\n")
        return 0
    print(__doc__)
    return 2


if __name__ == "__main__":
    sys.exit(main(sys.argv))
