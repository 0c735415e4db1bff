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
