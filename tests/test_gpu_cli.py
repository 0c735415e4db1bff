"""The `point_converter` command line (reference: point-converter/src/main.rs + lib.rs:11-60) end to end on the GPU:
PLY files in, node files + metadata.json out, compared with the oracle fed the same points."""
import struct
import subprocess

import numpy as np
import pytest

from pcutil import ROOT, compare_trees, make_points, oracle_convert, read_tree

pytestmark = pytest.mark.gpu
CLI = ROOT / "point-cloud_b200" / "point_converter"


def _write_ply(path, pts):
    hdr = ("ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty float x\nproperty float y\n"
           "property float z\nproperty uchar red\nproperty uchar green\nproperty uchar blue\nproperty uchar alpha\n"
           "end_header\n" % len(pts))
    with open(path, "wb") as f:
        f.write(hdr.encode())
        f.write(pts.tobytes())


def test_cli_files_dirs_and_merge(tmp_path):
    rng = np.random.default_rng(31)
    A = make_points(rng.random((60_000, 3)) * [2500, 1800, 60] - [1200, 900, 10])
    B = make_points(rng.random((25_000, 3)) * [2500, 1800, 60] - [1200, 900, 10])
    C = make_points(rng.random((12_345, 3)) * [300, 300, 30])
    (tmp_path / "in").mkdir()
    _write_ply(tmp_path / "a.ply", A)
    _write_ply(tmp_path / "in" / "b.ply", B)
    (tmp_path / "in" / "notes.txt").write_text("ignored")
    out_g, out_o = tmp_path / "gpu", tmp_path / "oracle"
    r = subprocess.run([str(CLI), "-o", str(out_g), "-f", str(tmp_path / "a.ply"), "-d", str(tmp_path / "in")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Unsupported file format 'txt'" in r.stderr and "Finished converting after" in r.stderr
    oracle_convert([A, B], out_dir=str(out_o))
    compare_trees(read_tree(out_o), read_tree(out_g), "oracle", "cli")
    assert (out_o / "metadata.json").read_text() == (out_g / "metadata.json").read_text()
    # second invocation on the same output directory: merge into the existing metadata.json (lib.rs:86-101)
    _write_ply(tmp_path / "c.ply", C)
    r = subprocess.run([str(CLI), "--output", str(out_g), "--files", str(tmp_path / "c.ply")], capture_output=True, text=True)
    assert r.returncode == 0 and "Found an existing metadata file." in r.stderr, r.stderr
    oracle_convert([C], out_dir=str(out_o))
    compare_trees(read_tree(out_o), read_tree(out_g), "oracle-merge", "cli-merge")
    assert (out_o / "metadata.json").read_text() == (out_g / "metadata.json").read_text()


def test_cli_reingest_own_format_conserves_points(tmp_path):
    rng = np.random.default_rng(32)
    A = make_points(rng.random((40_000, 3)) * [90, 90, 20])
    _write_ply(tmp_path / "a.ply", A)
    r = subprocess.run([str(CLI), "-o", str(tmp_path / "one"), "-f", str(tmp_path / "a.ply")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(CLI), "-o", str(tmp_path / "two"), "-f", str(tmp_path / "one" / "metadata.json")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    from pcutil import sort_records, tree_points
    assert np.array_equal(sort_records(tree_points(read_tree(tmp_path / "two"))), sort_records(A))


def test_cli_without_inputs_warns(tmp_path):
    r = subprocess.run([str(CLI), "-o", str(tmp_path / "x")], capture_output=True, text=True)
    assert r.returncode == 0 and "Please provide some files or directories" in r.stderr   # main.rs:40-43
    assert not (tmp_path / "x").exists()


def _write_las(path, ints, rgb16, fmt, rec_len, scale, offset, colour_off):
    """Minimal uncompressed LAS 1.2 file: 227-byte header + fixed-stride records (X, Y, Z as i32, colour as 3 x u16)."""
    n = len(ints)
    hdr = bytearray(227)
    hdr[0:4] = b"LASF"
    hdr[24], hdr[25] = 1, 2
    struct.pack_into("<H", hdr, 94, 227)
    struct.pack_into("<I", hdr, 96, 227)
    hdr[104] = fmt
    struct.pack_into("<H", hdr, 105, rec_len)
    struct.pack_into("<I", hdr, 107, n)
    struct.pack_into("<6d", hdr, 131, *scale, *offset)
    rec = np.zeros((n, rec_len), dtype=np.uint8)
    rec[:, 0:12] = np.ascontiguousarray(ints, dtype="<i4").view(np.uint8).reshape(n, 12)
    rec[:, 12:20] = 0x5A                                          # intensity, flags, ... : ignored by the reader
    if colour_off >= 0:
        rec[:, colour_off:colour_off + 6] = np.ascontiguousarray(rgb16, dtype="<u2").view(np.uint8).reshape(n, 6)
    with open(path, "wb") as f:
        f.write(bytes(hdr))
        f.write(rec.tobytes())


@pytest.mark.parametrize("fmt,rec_len,colour_off", [(2, 26, 20), (3, 34, 28), (1, 28, -1), (7, 37, 30)])
def test_cli_las_decoded_on_the_device(tmp_path, fmt, rec_len, colour_off):
    """A .las file through the GPU path: raw records are copied to the device and decoded by
    k_las_decode - f64 scale + offset then `as f32`, colour = low byte of each u16 channel, alpha 255
    (converter/las.rs:32-41).  The tree must equal the oracle's fed the points a host-side restatement of las.rs gives
    (numpy f64 arithmetic), and the CLI's host decode (--host-las) must give the same bytes.  Odd record lengths
    exercise the unaligned staging; 130 001 points = several decode blocks + a ragged tail, 13 batches + a short one."""
    rng = np.random.default_rng(100 + fmt)
    n = 130_001
    ints = np.stack([rng.integers(-150_000, 150_000, n), rng.integers(-120_000, 120_000, n), rng.integers(0, 9_000, n)], axis=1)
    ints[0] = (2_147_483_647, -2_147_483_648, 0)                  # extremes of the integer range
    ints[0] = ints[0] // 4096                                     # (kept inside the level-0 index range)
    rgb16 = rng.integers(0, 65536, (n, 3))
    scale, offset = (0.01, 0.01, 0.001), (100.0, -50.0, 0.5)
    _write_las(tmp_path / "scan.las", ints, rgb16, fmt, rec_len, scale, offset, colour_off)
    xyz = (ints.astype(np.float64) * np.array(scale) + np.array(offset)).astype(np.float32)
    rgba = np.zeros((n, 4), dtype=np.uint8)
    if colour_off >= 0:
        rgba[:, :3] = (rgb16 & 0xFF).astype(np.uint8)
    rgba[:, 3] = 255
    pts = make_points(xyz, rgba)
    out_g, out_h, out_o = tmp_path / "gpu", tmp_path / "host", tmp_path / "oracle"
    r = subprocess.run([str(CLI), "-o", str(out_g), "-f", str(tmp_path / "scan.las")], capture_output=True, text=True)
    assert r.returncode == 0 and f"Converting {n} points" in r.stderr, r.stderr
    r = subprocess.run([str(CLI), "-o", str(out_h), "--host-las", "-f", str(tmp_path / "scan.las")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    oracle_convert([pts], out_dir=str(out_o))
    tg, th = read_tree(out_g), read_tree(out_h)
    compare_trees(read_tree(out_o), tg, "oracle", "cli-las-device")
    assert tg == th, "device and host LAS decode differ"
    assert (out_o / "metadata.json").read_text() == (out_g / "metadata.json").read_text() == (out_h / "metadata.json").read_text()


def test_cli_streams_many_chunks_and_truncated_file(tmp_path):
    """The CLI reads its input in bounded page-locked chunks (lib.rs:31-52 holds one batch at a time).  A PLY larger
    than one chunk (4.19 M points) crosses chunk boundaries; a second, truncated PLY ends at its last complete batch
    (lib.rs:36-39).  Compared with the oracle fed the same points."""
    rng = np.random.default_rng(77)
    A = make_points(rng.random((4_400_000, 3)) * [1900, 1900, 80] - [950, 950, 10])
    B = make_points(rng.random((25_500, 3)) * [500, 500, 30])
    _write_ply(tmp_path / "a.ply", A)
    _write_ply(tmp_path / "b.ply", B)
    data = (tmp_path / "b.ply").read_bytes()
    (tmp_path / "b.ply").write_bytes(data[:-100])                 # the third batch of B is incomplete
    out_g, out_o = tmp_path / "gpu", tmp_path / "oracle"
    r = subprocess.run([str(CLI), "-o", str(out_g), "-f", str(tmp_path / "a.ply"), str(tmp_path / "b.ply")],
                       capture_output=True, text=True)
    assert r.returncode == 0 and "unexpected end of file" in r.stderr, r.stderr
    oracle_convert([A, B[:20_000]], out_dir=str(out_o))
    compare_trees(read_tree(out_o), read_tree(out_g), "oracle", "cli-streamed")
    assert (out_o / "metadata.json").read_text() == (out_g / "metadata.json").read_text()
