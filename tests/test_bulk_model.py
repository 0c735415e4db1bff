"""The bulk (level-synchronous) formulation used by the CUDA path equals the sequential oracle.
CPU only; validates SURVEY.md §8 N3 (arrival keys, emission order, flush rule, merge state)."""
import tempfile

import numpy as np
import pytest

from bulk_model import bulk_build
from pcutil import as_points, compare_trees, make_points, oracle_convert, read_tree


def _raw(a):
    return as_points(a).view(np.uint8).reshape(-1, 16)


def _check(files, config=None, bs=10_000):
    o = oracle_convert(files, config=config, batch_size=bs)
    nodes, meta = bulk_build([_raw(f) for f in files], config=config or (5000, 96, 1000.0), batch_size=bs)
    assert meta["number_of_points"] == o.number_of_points
    assert meta["hierarchies"] == o.hierarchies
    assert meta.get("arrivals", 0) == o.arrivals
    compare_trees(o.nodes, nodes, "oracle", "bulk")
    assert all(o.nodes[k] == nodes[k] for k in nodes)  # canonical writers agree byte for byte


def test_unit_cube_chain():
    rng = np.random.default_rng(1)
    _check([make_points(rng.random((30000, 3)))])


def test_small_config_and_ties():
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


@pytest.mark.parametrize("L", [1, 2, 3, 8, 16, 63, 64, 65])
def test_limit_edges(L):
    rng = np.random.default_rng(L)
    p = (0.3 + rng.random((64, 3)) * 0.05).astype(np.float32)
    for bs in (1, 2, 3, 7, 64, 100):
        _check([make_points(p)], config=(L, 4, 8.0), bs=bs)


def _merge(A, B, cfg, bs):
    with tempfile.TemporaryDirectory() as d:
        oracle_convert([A], out_dir=d, config=cfg, batch_size=bs)
        first = read_tree(d)
        o2 = oracle_convert([B], out_dir=d, batch_size=bs)
        final = read_tree(d)
    n1, m1 = bulk_build([_raw(A)], config=cfg, batch_size=bs)
    compare_trees(first, n1, "oracle1", "bulk1")
    n2, m2 = bulk_build([_raw(B)], config=cfg, batch_size=bs, existing=n1,
                        meta0=dict(number_of_points=m1["number_of_points"], hierarchies=m1["hierarchies"],
                                   bbox=m1["bbox"]))
    merged = dict(n1)
    merged.update(n2)
    compare_trees(final, merged, "oracle-merge", "bulk-merge")
    assert m2["hierarchies"] == o2.hierarchies and m2["number_of_points"] == o2.number_of_points
    assert np.array_equal(np.concatenate(m2["bbox"]), o2.bbox)


def test_merge_mode():
    rng = np.random.default_rng(7)
    _merge(make_points(rng.random((20000, 3)) * 40 - 20), make_points(rng.random((15000, 3)) * 50 - 25),
           (100, 8, 16.0), 1000)
    q = lambda n: np.round(rng.random((n, 3)) * 64) / 8 - 4
    _merge(make_points(q(9000)), make_points(q(9000)), (64, 6, 4.0), 512)
    for L in (3, 8, 16):
        for bs in (1, 5, 64):
            p = lambda: (0.3 + rng.random((40, 3)) * 0.05).astype(np.float32)
            _merge(make_points(p()), make_points(p()), (L, 4, 8.0), bs)
