"""Known-answer tests that pin the CPU oracle (oracle/pc_oracle*.cpp) to the reference's source.

The reference holds no tests or fixtures for this path and cannot be compiled here (Rust), so
every expected value below is derived BY HAND from the cited reference lines (paths relative to
/root/reference).  CPU only."""
import ctypes as C
import struct

import numpy as np
import pytest

from pcutil import Node, make_points, oracle_convert, oracle_lib

f32 = np.float32


def _cell_size(h, max_cell_size=1000.0, dim=96, limit=5000):
    return oracle_lib().orc_cell_size(limit, dim, C.c_float(max_cell_size), h)


def _cell_index(p, h, max_cell_size=1000.0):
    pa = np.asarray(p, dtype=np.float32)
    out = np.zeros(3, dtype=np.int32)
    oracle_lib().orc_cell_index(C.c_float(max_cell_size), h, pa.ctypes.data, out.ctypes.data)
    return tuple(int(v) for v in out)


def _hex_from_world(p, R):
    pa = np.asarray(p, dtype=np.float32)
    out = np.zeros(3, dtype=np.int32)
    oracle_lib().orc_hex_from_world(pa.ctypes.data, C.c_float(R), out.ctypes.data)
    return tuple(int(v) for v in out)


def _hex_to_world(k, R):
    ka = np.asarray(k, dtype=np.int32)
    out = np.zeros(3, dtype=np.float32)
    oracle_lib().orc_hex_to_world(ka.ctypes.data, C.c_float(R), out.ctypes.data)
    return out


# -- KAT 1: metadata.rs:91-93  cell_size(h) = max_cell_size / (2u32.pow(h) as f32)
def test_cell_size_is_exact_power_of_two_scaling():
    for h in range(0, 21):
        assert _cell_size(h) == f32(1000.0) / f32(2 ** h)
        assert float(_cell_size(h)) * 2 ** h == 1000.0  # exactly representable
    assert _cell_size(3, 333.3) == f32(333.3) / f32(8.0)


# -- KAT 2: metadata.rs:100-102  (pos / cell_size).floor().as_ivec3(); nesting
def test_cell_index_hand_values_and_nesting():
    assert _cell_index((-0.5, 0.0, 999.9999), 0) == (-1, 0, 0)
    assert _cell_index((1000.0, -1000.0, -1000.0001), 0) == (1, -1, -2)
    assert _cell_index((499.99997, 500.0, -0.0), 1) == (0, 1, 0)      # -0.0 / 500 = -0.0, floor -> -0 -> 0
    rng = np.random.default_rng(0)
    pts = (rng.random((2000, 3)) * 10000 - 5000).astype(np.float32)
    for p in pts[:300]:
        prev = _cell_index(p, 0)
        for h in range(1, 16):
            cur = _cell_index(p, h)
            assert tuple(c >> 1 for c in cur) == prev  # arithmetic shift, negatives included
            prev = cur


def test_cell_index_saturates_like_rust_as_i32():
    # Rust `f32 as i32` saturates and maps NaN to 0
    assert _cell_index((3.0e12, -3.0e12, float("nan")), 0) == (2 ** 31 - 1, -(2 ** 31), 0)


# -- KAT 3: metadata.rs:104-106  index.as_vec3() * cell_size + cell_size / 2.0
def test_cell_pos():
    idx = np.array([-1, 0, 2], dtype=np.int32)
    out = np.zeros(3, dtype=np.float32)
    oracle_lib().orc_cell_pos(idx.ctypes.data, C.c_float(500.0), out.ctypes.data)
    assert tuple(out) == (-250.0, 250.0, 1250.0)


# -- KAT 4: hex.rs:17-85
def test_hex_offset_axial_roundtrip_and_centres():
    R = f32(0.5)
    for k in [(0, 0, 0), (3, 2, 1), (-4, -3, -2), (5, -7, 9), (-6, 5, -1), (10, 11, 0), (-10, -11, 3)]:
        w = _hex_to_world(k, R)
        # a point just inside the prism (centre in xy, slightly above the reference face in |z|)
        z = w[2] + (f32(0.25) * R if k[2] >= 0 else -f32(0.25) * R)
        assert _hex_from_world((w[0], w[1], z), R) == k


def test_hex_z_truncates_toward_zero():
    R = f32(1.0)
    # hex.rs:83  h = (pos.z / cell_radius) as i32  — truncation, so layer 0 is (-R, R)
    assert _hex_from_world((0.0, 0.0, -0.3), R)[2] == 0
    assert _hex_from_world((0.0, 0.0, 0.99), R)[2] == 0
    assert _hex_from_world((0.0, 0.0, -0.99), R)[2] == 0
    assert _hex_from_world((0.0, 0.0, -1.2), R)[2] == -1
    assert _hex_from_world((0.0, 0.0, 1.2), R)[2] == 1


def test_hex_to_world_formula():
    # hex.rs:18-24 then 55-65 with R = 2: offset (3, 2, 5) -> axial q = 3 - (2 - 0)/2 = 2, r = 2
    R = f32(2.0)
    S3 = f32(1.73205080757)
    w = _hex_to_world((3, 2, 5), R)
    x = R * (S3 * f32(2.0) + S3 / f32(2.0) * f32(2.0))
    y = R * f32(3.0) / f32(2.0) * f32(2.0)
    assert w[0] == x and w[1] == y and w[2] == f32(10.0)
    # odd row: offset (3, 3, 0) -> q = 3 - (3 - 1)/2 = 2, r = 3
    w = _hex_to_world((3, 3, 0), R)
    assert w[0] == R * (S3 * f32(2.0) + S3 / f32(2.0) * f32(3.0))
    # negative odd row: y = -3 -> (-3 & 1) = 1 -> (-3 - 1) / 2 = -2 -> q = x + 2
    w = _hex_to_world((0, -3, 0), R)
    assert w[0] == R * (S3 * f32(2.0) + S3 / f32(2.0) * f32(-3.0))


# ---- conversions on tiny clouds ---------------------------------------------------------------------
CFG = (5, 4, 8.0)  # limit 5, sub_grid_dimension 4, max_cell_size 8  =>  level-0 hex radius R = 8/4/2 = 1


def _run(xyz, bs, config=CFG):
    o = oracle_convert([make_points(np.asarray(xyz, dtype=np.float32))], config=config, batch_size=bs)
    return o, {k: Node(v) for k, v in o.nodes.items()}


def _xyz(node_rows):
    return [tuple(np.frombuffer(r[:12].tobytes(), dtype=np.float32)) for r in node_rows]


# -- KAT 5: cell.rs:70-94 — nearer point keeps the prism whatever the order; equal d2 keeps the earlier one
def test_contest_nearest_wins_and_ties_keep_first():
    # hex (0,0,0) with R = 1: reference centre (0,0,0); three points in that prism
    a, b, c = (0.10, 0.0, 0.5), (0.05, 0.0, 0.5), (0.4, 0.0, 0.5)
    for order in ([a, b, c], [c, b, a], [b, c, a]):
        _, nodes = _run(order, bs=10)
        root = nodes["h_0/c_0_0_0.bin"]
        assert _xyz(root.grid) == [tuple(np.float32(v) for v in b)]
        assert root.header["total"] == 3 and root.header["npoints"] == 1 and root.header["noverflow"] == 2
    # equal distance: mirror images about the centre of prism (1,0,0), cx = R * sqrt(3) (hex.rs:55-65); both
    # cx + 0.25 and cx - 0.25 are exact in f32, so the two d2 are bit-equal and the first arrival stays
    # (strict `<`, cell.rs:80)
    cx = f32(1.0) * (f32(1.73205080757) * f32(1.0))
    p, q = (cx + f32(0.25), 0.1, 0.5), (cx - f32(0.25), 0.1, 0.5)
    assert _hex_from_world(p, 1.0) == _hex_from_world(q, 1.0) == (1, 0, 0)
    for first, second in ((p, q), (q, p)):
        _, nodes = _run([first, second], bs=10)
        assert set(nodes) == {"h_0/c_0_0_0.bin"}
        assert _xyz(nodes["h_0/c_0_0_0.bin"].grid) == [tuple(np.float32(v) for v in first)]


def test_contest_emission_order_is_trigger_order():
    # arrivals in one prism with decreasing distance: each newcomer displaces the incumbent, which is emitted
    # at that moment (cell.rs:80-83); the overflow list therefore holds them in arrival order
    pts = [(0.4, 0.0, 0.5), (0.3, 0.0, 0.5), (0.2, 0.0, 0.5), (0.1, 0.0, 0.5)]
    _, nodes = _run(pts, bs=10)
    root = nodes["h_0/c_0_0_0.bin"]
    assert _xyz(root.grid) == [tuple(np.float32(v) for v in pts[3])]
    (child, lst), = root.entries.items()
    assert child == (0, 0, 0)
    assert _xyz(lst) == [tuple(np.float32(v) for v in p) for p in pts[:3]]


# -- KAT 6: cell.rs:108-153 — the overflow state machine, `<=` on first fill (:118) vs `<` later (:137)
def _same_prism(n):
    # n distinct points of hex prism (0,0,0) of level 0 (R = 1), all in child octant (0,0,0) of cell (0,0,0),
    # with strictly increasing distance so the first one stays the winner and losers leave in arrival order
    return [(0.01 + 0.001 * i, 0.0, 0.5) for i in range(n)]


def test_overflow_exactly_limit_in_one_call_is_kept():
    _, nodes = _run(_same_prism(6), bs=10)           # 1 winner + 5 losers in ONE call: 5 <= 5 -> Some
    root = nodes["h_0/c_0_0_0.bin"]
    assert set(nodes) == {"h_0/c_0_0_0.bin"}
    assert root.header["noverflow"] == 5 and root.header["total"] == 6
    assert len(root.entries[(0, 0, 0)]) == 5


def test_overflow_limit_reached_over_two_calls_is_flushed():
    o, nodes = _run(_same_prism(6), bs=5)            # call 1: 4 losers (kept); call 2: +1 -> 5 !< 5 -> flushed
    root = nodes["h_0/c_0_0_0.bin"]
    assert root.entries[(0, 0, 0)] is None           # marker None is written as count 0 (cell.rs:169-177)
    assert root.header["noverflow"] == 0 and root.header["total"] == 1
    assert "h_1/c_0_0_0.bin" in nodes and o.hierarchies >= 2
    assert sum(int(n.header["total"]) for n in nodes.values()) == 6


def test_overflow_more_than_limit_in_one_call_goes_straight_down():
    _, nodes = _run(_same_prism(7), bs=10)           # 6 losers in one call: 6 > 5 -> None, header untouched
    root = nodes["h_0/c_0_0_0.bin"]
    assert root.entries[(0, 0, 0)] is None
    assert root.header["noverflow"] == 0 and root.header["total"] == 1
    assert int(nodes["h_1/c_0_0_0.bin"].header["total"]) + sum(
        int(n.header["total"]) for k, n in nodes.items() if k.startswith("h_") and not k.startswith(("h_0", "h_1"))) == 6


def test_after_none_everything_passes_through():
    _, nodes = _run(_same_prism(7) + [(0.5, 0.0, 0.5)], bs=7)   # second call: 1 more loser, state None
    root = nodes["h_0/c_0_0_0.bin"]
    assert root.entries[(0, 0, 0)] is None and root.header["total"] == 1
    assert sum(int(n.header["total"]) for n in nodes.values()) == 8


# -- KAT 7: header accounting invariant (cell.rs:89-90, 119-120, 138-142)
def test_header_totals_add_up():
    rng = np.random.default_rng(5)
    o, nodes = _run(rng.random((3000, 3)) * 7.9, bs=97, config=(20, 4, 8.0))
    stored = 0
    for n in nodes.values():
        kept = sum(len(v) for v in n.entries.values() if v is not None)
        assert n.header["npoints"] == len(n.grid) and n.header["noverflow"] == kept
        assert n.header["total"] == n.header["npoints"] + n.header["noverflow"]
        stored += int(n.header["total"])
    assert stored == 3000 == o.number_of_points


# -- KAT 8: codec (cell.rs:155-229, 280-335; point.rs:25-37): bytes assembled by hand
def test_node_file_layout_bytes():
    pts = [(0.05, 0.0, 0.5), (0.4, 0.0, 0.5)]
    o, _ = _run(pts, bs=10)
    raw = o.nodes["h_0/c_0_0_0.bin"]
    expect = struct.pack("<I3i3I", 0, 0, 0, 0, 2, 1, 1)                   # hierarchy, index, total, grid, overflow
    expect += struct.pack("<2f3f", 8.0, 2.0, 4.0, 4.0, 4.0)               # size, sub_cell_size = 8/4, centre
    expect += struct.pack("<3f", *pts[0]) + bytes([0, 0, 0, 0])           # grid point (colour = ordinal 0)
    expect += struct.pack("<B", 1)                                        # one overflow entry
    expect += struct.pack("<3iI", 0, 0, 0, 1)                             # child (0,0,0), 1 point
    expect += struct.pack("<3f", *pts[1]) + bytes([1, 0, 0, 0])
    assert raw == expect


def test_metadata_json_text():
    o, _ = _run([(0.5, 0.25, 1e-7), (-1.5, 2.0, 3.0)], bs=10, config=(5000, 96, 1000.0))
    assert o.meta_json == (
        '{\n  "version": "1.0",\n  "name": "Unknown",\n  "number_of_points": 2,\n  "hierarchies": 1,\n'
        '  "bounding_box": {\n    "min": [\n      -1.5,\n      0.25,\n      1e-7\n    ],\n'
        '    "max": [\n      0.5,\n      2.0,\n      3.0\n    ]\n  },\n'
        '  "config": {\n    "cell_point_overflow_limit": 5000,\n    "sub_grid_dimension": 96,\n'
        '    "max_cell_size": 1000.0\n  }\n}')


# -- KAT 9: degenerate chain of single cells for a small-extent cloud
def test_unit_cube_chain_of_single_cells():
    rng = np.random.default_rng(1)
    o = oracle_convert([make_points(rng.random((25_000, 3)))])
    for h in range(0, 8):
        level = [k for k in o.nodes if k.startswith(f"h_{h}/")]
        assert level == [f"h_{h}/c_0_0_0.bin"]
    assert o.hierarchies >= 9


# -- KAT 10: merge == continued stream when batch boundaries coincide
def test_merge_equals_continued_stream(tmp_path):
    from pcutil import compare_trees, read_tree
    rng = np.random.default_rng(3)
    A = make_points(rng.random((4000, 3)) * 30 - 15)
    B = make_points(rng.random((3500, 3)) * 30 - 15)
    cfg, bs = (50, 6, 16.0), 500
    d1, d2 = tmp_path / "two_runs", tmp_path / "one_run"
    oracle_convert([A], out_dir=str(d1), config=cfg, batch_size=bs)
    oracle_convert([B], out_dir=str(d1), batch_size=bs)
    oracle_convert([A, B], out_dir=str(d2), config=cfg, batch_size=bs)
    compare_trees(read_tree(d1), read_tree(d2), "merge", "single")
    assert (d1 / "metadata.json").read_text() == (d2 / "metadata.json").read_text()


def test_empty_batch_still_visits_level_zero():
    # converter.rs:111,120: add_points_in_hierarchy(0, ..) runs create_hierarchy_folder(0) unconditionally
    o = oracle_convert([make_points(np.zeros((0, 3)))])
    assert o.hierarchies == 1 and o.number_of_points == 0 and not o.nodes


def test_lru_residency_is_transparent(tmp_path):
    """converter.rs:92, 160-178: the reference keeps at most 100 cells in RAM and spills evicted cells as whole node
    files.  The oracle's emulation of that (used by bench.py's "what the real CLI pays" CPU line) must not change a
    single byte of the result: a capacity of 3 cells against everything-in-RAM, fresh run and merge run."""
    from pcutil import compare_trees, read_tree
    rng = np.random.default_rng(5)
    a = make_points(rng.random((30_000, 3)) * 90 - 45)
    b = make_points(rng.random((20_000, 3)) * 120 - 60)
    cfg = (60, 8, 32.0)
    d0, d1 = str(tmp_path / "ram"), str(tmp_path / "lru")
    oracle_convert([a], out_dir=d0, config=cfg, batch_size=1000, dump=False)
    oracle_convert([b], out_dir=d0, batch_size=1000, dump=False)
    oracle_convert([a], out_dir=d1, config=cfg, batch_size=1000, dump=False, lru=3)
    oracle_convert([b], out_dir=d1, batch_size=1000, dump=False, lru=3)
    t0, t1 = read_tree(d0), read_tree(d1)
    assert len(t0) > 20
    compare_trees(t0, t1, "in-ram", "lru-3")
    assert open(d0 + "/metadata.json").read() == open(d1 + "/metadata.json").read()
