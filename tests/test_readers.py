"""Input readers of the CLI front end (point-cloud_b200/csrc/reader.cpp) — host logic, CPU only.
Reference behaviour: converter/ply.rs:35-82 + point.rs:56-131, converter/las.rs:22-55, converter/own.rs:15-87."""
import struct
import subprocess

import numpy as np
import pytest

from pcutil import POINT_DTYPE, ROOT, make_points, oracle_convert

CLI = ROOT / "point-cloud_b200" / "point_converter"


def _read(paths, tmp_path):
    out = tmp_path / "dump.bin"
    r = subprocess.run([str(CLI), "--dump-points", str(out), "-f", *map(str, paths)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return np.fromfile(out, dtype=POINT_DTYPE), r.stderr


def _ply(path, header_props, rows, fmt="binary_little_endian", endian="<", rowfmt=""):
    hdr = "ply\nformat %s 1.0\ncomment test\nelement vertex %d\n" % (fmt, len(rows))
    hdr += "".join("property %s %s\n" % p for p in header_props) + "end_header\n"
    with open(path, "wb") as f:
        f.write(hdr.encode())
        if fmt == "ascii":
            for r in rows:
                f.write((" ".join(str(v) for v in r) + "\n").encode())
        else:
            for r in rows:
                f.write(struct.pack(endian + rowfmt, *r))


def test_ply_binary_le_float_uchar(tmp_path):
    rows = [(1.5, -2.25, 3.0, 10, 20, 30), (0.1, 0.2, 0.3, 255, 0, 7)]
    _ply(tmp_path / "a.ply", [("float", "x"), ("float", "y"), ("float", "z"), ("uchar", "red"), ("uchar", "green"),
                              ("uchar", "blue")], rows, rowfmt="3f3B")
    pts, _ = _read([tmp_path / "a.ply"], tmp_path)
    assert len(pts) == 2
    assert (pts["x"][0], pts["y"][0], pts["z"][0]) == (np.float32(1.5), np.float32(-2.25), np.float32(3.0))
    assert pts["rgba"].tolist() == [[10, 20, 30, 255], [255, 0, 7, 255]]   # alpha defaults to 255 (point.rs:16-23)


def test_ply_big_endian_double_and_float_colour_quirk(tmp_path):
    # doubles are narrowed `as f32` (point.rs:69-71); float colours are divided by 255 then truncated (point.rs:96-98)
    rows = [(0.1, 1e10, -7.0, 255.0, 254.9, 1000000.0, 3)]
    _ply(tmp_path / "b.ply", [("double", "x"), ("double", "y"), ("double", "z"), ("float", "r"), ("float", "g"),
                              ("float", "b"), ("uchar", "a")], rows, fmt="binary_big_endian", endian=">", rowfmt="3d3fB")
    pts, _ = _read([tmp_path / "b.ply"], tmp_path)
    assert pts["x"][0] == np.float32(0.1) and pts["y"][0] == np.float32(1e10)
    assert pts["rgba"].tolist() == [[1, 0, 255, 3]]


def test_ply_other_coordinate_types_are_ignored(tmp_path):
    _ply(tmp_path / "c.ply", [("int", "x"), ("float", "y"), ("float", "z"), ("short", "extra")], [(5, 1.0, 2.0, 9)],
         rowfmt="i2fh")
    pts, _ = _read([tmp_path / "c.ply"], tmp_path)
    assert (pts["x"][0], pts["y"][0], pts["z"][0]) == (0.0, 1.0, 2.0)     # point.rs:74 `_ => {}`


def test_ply_ascii_delivers_no_points(tmp_path):
    _ply(tmp_path / "d.ply", [("float", "x"), ("float", "y"), ("float", "z")], [(1, 2, 3), (4, 5, 6)], fmt="ascii")
    pts, log = _read([tmp_path / "d.ply"], tmp_path)
    assert len(pts) == 0 and "total 2 read 0 ascii 1" in log               # ply.rs:43-51


def test_ply_truncated_file(tmp_path):
    rows = [(float(i), 0.0, 0.0) for i in range(10)]
    _ply(tmp_path / "e.ply", [("float", "x"), ("float", "y"), ("float", "z")], rows, rowfmt="3f")
    data = (tmp_path / "e.ply").read_bytes()
    (tmp_path / "e.ply").write_bytes(data[:-30])
    pts, log = _read([tmp_path / "e.ply"], tmp_path)
    assert len(pts) == 7 and "status 2" in log


def test_las_uncompressed_format2(tmp_path):
    n = 3
    hdr = bytearray(227)
    hdr[0:4] = b"LASF"
    hdr[24], hdr[25] = 1, 2
    struct.pack_into("<H", hdr, 94, 227)
    struct.pack_into("<I", hdr, 96, 227)
    hdr[104] = 2
    struct.pack_into("<H", hdr, 105, 26)
    struct.pack_into("<I", hdr, 107, n)
    struct.pack_into("<6d", hdr, 131, 0.01, 0.01, 0.001, 100.0, -50.0, 0.5)
    recs = b""
    ints = [(12345, -200, 7), (0, 0, 0), (-1, 2147483647, -2147483648)]
    for X, Y, Z in ints:
        recs += struct.pack("<3iHBBbBH3H", X, Y, Z, 0, 0, 0, 0, 0, 0, 0x1234, 0xFF01, 0x00FE)
    (tmp_path / "f.las").write_bytes(bytes(hdr) + recs)
    pts, _ = _read([tmp_path / "f.las"], tmp_path)
    assert len(pts) == n
    for p, (X, Y, Z) in zip(pts, ints):
        assert p["x"] == np.float32(X * 0.01 + 100.0) and p["y"] == np.float32(Y * 0.01 - 50.0)
        assert p["z"] == np.float32(Z * 0.001 + 0.5)
        assert p["rgba"].tolist() == [0x34, 0x01, 0xFE, 255]               # `u16 as u8` keeps the low byte (las.rs:40)


def test_own_format_reingest_conserves_points(tmp_path):
    rng = np.random.default_rng(4)
    pts = make_points(rng.random((5000, 3)) * 20 - 10)
    oracle_convert([pts], out_dir=str(tmp_path / "cloud"), config=(50, 6, 16.0), batch_size=500)
    got, _ = _read([tmp_path / "cloud" / "metadata.json"], tmp_path)
    a = np.sort(got.view(np.uint8).reshape(-1, 16).view("V16").reshape(-1))
    b = np.sort(pts.view(np.uint8).reshape(-1, 16).view("V16").reshape(-1))
    assert len(got) == 5000 and np.array_equal(a, b)                       # own.rs:21-49: all_points of every cell


def test_unsupported_extension_is_skipped(tmp_path):
    (tmp_path / "x.txt").write_text("hello")
    pts, log = _read([tmp_path / "x.txt"], tmp_path)
    assert len(pts) == 0 and "Unsupported file format 'txt'" in log        # lib.rs:78-81
