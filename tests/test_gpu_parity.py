"""GPU parity: the CUDA path (through the C ABI, include/pcb.h) against the CPU oracle, bit for bit.
Integer/byte work: the bar is exact equality of every node image (canonical in-file order)."""
import tempfile

import numpy as np
import pytest

from pcutil import (as_points, compare_trees, load_pcb, make_points, oracle_convert, read_tree, sort_records,
                    tree_points)

pytestmark = pytest.mark.gpu


def _gpu(files, config=None, bs=0, out_dir=None):
    pcb = load_pcb()
    return pcb.convert([as_points(f) for f in files], out_dir, config=config, batch_size=bs)


def _check(files, config=None, bs=10_000):
    o = oracle_convert(files, config=config, batch_size=bs)
    nodes, mj, st, _ = _gpu(files, config=config, bs=bs)
    assert st.number_of_points == o.number_of_points
    assert st.hierarchies == o.hierarchies
    assert st.arrivals == o.arrivals
    compare_trees(o.nodes, nodes, "oracle", "gpu")
    assert all(o.nodes[k] == nodes[k] for k in nodes), "canonical node images differ byte-wise"
    assert mj == o.meta_json
    return st


def test_unit_cube_chain():
    rng = np.random.default_rng(1)
    _check([make_points(rng.random((30000, 3)))])


def test_small_config_ties_and_duplicates():
    rng = np.random.default_rng(2)
    _check([make_points(rng.random((30000, 3)) * 50 - 25)], config=(100, 8, 10.0), bs=1000)
    base = rng.random((3000, 3)).astype(np.float32) * 20 - 10
    _check([make_points(np.concatenate([base, base, base[::-1], base]))], config=(50, 4, 8.0), bs=700)
    q = np.round(rng.random((20000, 3)) * 64) / 8 - 4
    _check([make_points(q)], config=(64, 6, 4.0), bs=512)


def test_multi_file_with_empty_file():
    rng = np.random.default_rng(3)
    _check([make_points(rng.random((2500, 3)) * 30), make_points(rng.random((0, 3))),
            make_points(rng.random((1234, 3)) * 30 - 10)], config=(40, 5, 16.0), bs=300)


def test_only_empty_files():
    _check([make_points(np.zeros((0, 3)))])


@pytest.mark.parametrize("L", [1, 2, 3, 8, 16, 63, 64, 65])
def test_limit_edges(L):
    rng = np.random.default_rng(L)
    p = (0.3 + rng.random((64, 3)) * 0.05).astype(np.float32)
    for bs in (1, 2, 3, 7, 64, 100):
        _check([make_points(p)], config=(L, 4, 8.0), bs=bs)


def test_negative_cells_and_many_roots():
    rng = np.random.default_rng(5)
    _check([make_points(rng.random((200_000, 3)) * [4000, 4000, 300] - [2000, 2000, 100])])


def test_default_config_1m_uniform():
    """BASELINE.json configs[0]: 1M uniform points in a unit cube, default node capacity."""
    rng = np.random.default_rng(11)
    st = _check([make_points(rng.random((1_000_000, 3)))])
    assert st.gpu_launches > 0


def test_files_on_disk_and_merge():
    rng = np.random.default_rng(7)
    A = make_points(rng.random((20000, 3)) * 40 - 20)
    B = make_points(rng.random((15000, 3)) * 50 - 25)
    cfg, bs = (100, 8, 16.0), 1000
    with tempfile.TemporaryDirectory() as d_o, tempfile.TemporaryDirectory() as d_g:
        oracle_convert([A], out_dir=d_o, config=cfg, batch_size=bs)
        _gpu([A], config=cfg, bs=bs, out_dir=d_g)
        compare_trees(read_tree(d_o), read_tree(d_g), "oracle1", "gpu1")
        assert open(f"{d_o}/metadata.json").read() == open(f"{d_g}/metadata.json").read()
        oracle_convert([B], out_dir=d_o, batch_size=bs)
        _gpu([B], bs=bs, out_dir=d_g)
        compare_trees(read_tree(d_o), read_tree(d_g), "oracle-merge", "gpu-merge")
        assert open(f"{d_o}/metadata.json").read() == open(f"{d_g}/metadata.json").read()


def test_merge_limit_edges():
    rng = np.random.default_rng(9)
    for L in (3, 8, 16):
        for bs in (1, 5, 64):
            A = make_points((0.3 + rng.random((40, 3)) * 0.05).astype(np.float32))
            B = make_points((0.3 + rng.random((40, 3)) * 0.05).astype(np.float32))
            with tempfile.TemporaryDirectory() as d_o, tempfile.TemporaryDirectory() as d_g:
                oracle_convert([A], out_dir=d_o, config=(L, 4, 8.0), batch_size=bs)
                oracle_convert([B], out_dir=d_o, batch_size=bs)
                _gpu([A], config=(L, 4, 8.0), bs=bs, out_dir=d_g)
                _gpu([B], bs=bs, out_dir=d_g)
                compare_trees(read_tree(d_o), read_tree(d_g), "oracle-merge", "gpu-merge")


def test_point_conservation_large():
    """Size-independent property at a size the oracle does not need to run: every input point is stored
    exactly once (grid or overflow), header totals add up."""
    rng = np.random.default_rng(13)
    pts = make_points(rng.random((3_000_000, 3)) * [3000, 3000, 50] - [1500, 1500, 0])
    nodes, _, st, _ = _gpu([pts])
    stored = tree_points(nodes)
    assert len(stored) == len(pts) == st.number_of_points
    assert np.array_equal(sort_records(stored), sort_records(pts))
