"""The C++ oracle against the second, independent restatement (oracle/model.py, pure Python + np.float32
scalars written directly from the Rust sources): node images must agree bit for bit.  CPU only, small inputs."""
import sys
from pathlib import Path

import numpy as np
import pytest

from pcutil import ROOT, compare_trees, make_points, oracle_convert

sys.path.insert(0, str(ROOT / "oracle"))
import model  # noqa: E402


def _check(xyz, cfg, bs):
    pts = make_points(np.asarray(xyz, dtype=np.float32))
    o = oracle_convert([pts], config=cfg, batch_size=bs)
    m = model.Converter(model.Config(*cfg), batch_size=bs)
    m.add_file(pts)
    nodes = m.nodes()
    compare_trees(o.nodes, nodes, "oracle", "model")
    assert all(o.nodes[k] == nodes[k] for k in nodes)
    assert m.hierarchies == o.hierarchies and m.number_of_points == o.number_of_points
    assert np.array_equal(np.concatenate(m.bbox), o.bbox)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_clouds(seed):
    rng = np.random.default_rng(seed)
    _check(rng.random((1500, 3)) * 40 - 20, (30, 5, 16.0), 128)


def test_duplicates_and_quantised_coordinates():
    rng = np.random.default_rng(7)
    base = np.round(rng.random((400, 3)) * 32) / 4 - 4
    _check(np.concatenate([base, base[::-1], base]), (16, 4, 4.0), 100)


@pytest.mark.parametrize("L,bs", [(1, 1), (2, 3), (5, 5), (8, 64)])
def test_limit_edges(L, bs):
    rng = np.random.default_rng(L * 31 + bs)
    _check(0.3 + rng.random((80, 3)) * 0.05, (L, 4, 8.0), bs)
