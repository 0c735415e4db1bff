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


def test_clustered_blobs_skewed_tree():
    """BASELINE.json configs[3] in miniature: Zipf-weighted Gaussian blobs (deep, skewed subtrees)."""
    rng = np.random.default_rng(17)
    centres = rng.random((12, 3)) * [800, 800, 50] - [400, 400, 0]
    sigma = 100.0 ** rng.random(12) * 0.05
    w = 1.0 / np.arange(1, 13)
    k = rng.choice(12, size=400_000, p=w / w.sum())
    xyz = centres[k] + rng.standard_normal((400_000, 3)) * sigma[k, None]
    st = _check([make_points(xyz)])
    assert st.hierarchies >= 8


def test_terrain_like_default_config():
    """BASELINE.json configs[1] in miniature: 2.5-D surface over several level-0 cells, default config."""
    rng = np.random.default_rng(18)
    xy = rng.random((500_000, 2)) * 900 - 450
    z = 30 * np.sin(xy[:, 0] / 80) * np.cos(xy[:, 1] / 60) + rng.standard_normal(500_000) * 0.05
    _check([make_points(np.column_stack([xy, z]))])


def test_city_scan_in_miniature():
    """BASELINE.json configs[4] in miniature (synth.city: ground + box faces + sensor noise): a 300 m block with a coarse
    config, and 800 m of city with the default grid and a limit of 500."""
    import sys
    sys.path.insert(0, str(__import__("pathlib").Path(__file__).resolve().parent.parent / "point-cloud_b200"))
    import synth
    _check([synth.city(300_000, 5, "cpu", extent=300.0, boxes=40).numpy()], config=(50, 16, 64.0), bs=5000)
    _check([synth.city(600_000, 6, "cpu", extent=800.0, boxes=150).numpy()], config=(500, 96, 1000.0))


def test_payload_sort_path_forced():
    """The (key, position) pairs fallback of the level sort (taken when cell + hex + position bits exceed 64)
    must give the same trees; forced through PCB_FORCE_PAIRS in a fresh process."""
    import os
    import subprocess
    import sys
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import numpy as np\n"
        "from pcutil import compare_trees, load_pcb, make_points, oracle_convert\n"
        "rng = np.random.default_rng(19)\n"
        "for cfg, bs, n, ext in (((100, 8, 10.0), 1000, 40000, 50), ((5000, 96, 1000.0), 10000, 200000, 2500)):\n"
        "    pts = make_points(rng.random((n, 3)) * ext - ext / 2)\n"
        "    o = oracle_convert([pts], config=cfg, batch_size=bs)\n"
        "    nodes, mj, st, _ = load_pcb().convert([pts], config=cfg, batch_size=bs)\n"
        "    compare_trees(o.nodes, nodes, 'oracle', 'gpu-pairs')\n"
        "    assert mj == o.meta_json and all(o.nodes[k] == nodes[k] for k in nodes)\n"
        "print('pairs ok')\n") % os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, PCB_ENABLE_TUNING="1", PCB_FORCE_PAIRS="1")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "pairs ok" in r.stdout, r.stdout + r.stderr


def test_segmented_sort_forced():
    """Levels whose cell bits would cost extra radix passes sort runs of consecutive cells ("segments")
    independently (radix.cuh).  Forced on small inputs with PCB_SEG_MIN=1 (fresh process): many cells per level
    (several segments, partly filled tiles, segments of one tile), the default grid and a coarse one, with and
    without the payload sort."""
    import os
    import subprocess
    import sys
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import numpy as np\n"
        "from pcutil import compare_trees, load_pcb, make_points, oracle_convert\n"
        "rng = np.random.default_rng(43)\n"
        "for cfg, bs, n, ext in (((100, 8, 10.0), 1000, 80000, 90), ((5000, 96, 1000.0), 10000, 300000, 9000),\n"
        "                        ((20, 96, 4.0), 5000, 120000, 60), ((300, 20, 16.0), 2000, 250000, 100)):\n"
        "    pts = make_points(rng.random((n, 3)) * [ext, ext, ext / 4] - [ext / 2, ext / 2, ext / 8])\n"
        "    o = oracle_convert([pts], config=cfg, batch_size=bs)\n"
        "    nodes, mj, st, _ = load_pcb().convert([pts], config=cfg, batch_size=bs)\n"
        "    assert st.arrivals == o.arrivals and st.hierarchies == o.hierarchies\n"
        "    compare_trees(o.nodes, nodes, 'oracle', 'gpu-seg')\n"
        "    assert mj == o.meta_json and all(o.nodes[k] == nodes[k] for k in nodes)\n"
        "print('seg ok')\n") % os.path.dirname(os.path.abspath(__file__))
    for extra in ({}, {"PCB_FORCE_PAIRS": "1"}):
        env = dict(os.environ, PCB_ENABLE_TUNING="1", PCB_SEG_MIN="1", **extra)
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and "seg ok" in r.stdout, r.stdout + r.stderr


def test_record_filter_forced():
    """Record filter (level.cuh): a level is built from the early arrivals of every cell, a hash-table filter over
    the later ones and a second contest over early winners + candidates.  Forced on every level of small inputs
    (PCB_FILTER_SHIFT = early share 1/2, 1/4, 1/16; fresh processes): duplicate- and tie-heavy clouds, many cells,
    tiny cells (early share of one arrival), multi-file runs, with and without forced segmentation."""
    import os
    import subprocess
    import sys
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import numpy as np\n"
        "from pcutil import compare_trees, load_pcb, make_points, oracle_convert\n"
        "rng = np.random.default_rng(47)\n"
        "clouds = []\n"
        "for cfg, bs, n, ext in (((100, 8, 10.0), 1000, 80000, 30), ((5000, 96, 1000.0), 10000, 300000, 3000),\n"
        "                        ((20, 96, 4.0), 5000, 120000, 20), ((300, 20, 16.0), 2000, 250000, 100)):\n"
        "    clouds.append((cfg, bs, [make_points(rng.random((n, 3)) * [ext, ext, ext / 4] - [ext / 2, ext / 2, ext / 8])]))\n"
        "dup = np.round(rng.random((20000, 3)) * 12, 0) / 2\n"   # few distinct positions: ties and exact duplicates
        "clouds.append(((50, 8, 8.0), 700, [make_points(dup), make_points(dup[::-1].copy()), make_points(rng.random((5000, 3)) * 6)]))\n"
        "for cfg, bs, files in clouds:\n"
        "    o = oracle_convert(files, config=cfg, batch_size=bs)\n"
        "    nodes, mj, st, _ = load_pcb().convert(files, config=cfg, batch_size=bs)\n"
        "    assert st.arrivals == o.arrivals and st.hierarchies == o.hierarchies\n"
        "    compare_trees(o.nodes, nodes, 'oracle', 'gpu-filter')\n"
        "    assert mj == o.meta_json and all(o.nodes[k] == nodes[k] for k in nodes)\n"
        "print('filter ok')\n") % os.path.dirname(os.path.abspath(__file__))
    # the last one gives the filter up after round 1 (too many predicted candidates) and falls back to the full sort
    for extra in ({"PCB_FILTER_SHIFT": "1"}, {"PCB_FILTER_SHIFT": "2", "PCB_SEG_MIN": "1"}, {"PCB_FILTER_SHIFT": "4"},
                  {"PCB_FILTER_SHIFT": "3", "PCB_FILTER_BAIL": "1"}):
        env = dict(os.environ, PCB_ENABLE_TUNING="1", **extra)
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and "filter ok" in r.stdout, str(extra) + r.stdout[-2000:] + r.stderr[-3000:]


def test_waves_of_level0_cells_forced():
    """The forest is built a group of level-0 cells ("wave") at a time.  Tiny waves (PCB_WAVE_POINTS, fresh
    process) must give the same trees as the oracle: many roots incl. negative cells, the default config, and a
    merge into an existing output whose cells are spread over several waves."""
    import os
    import subprocess
    import sys
    code = (
        "import sys, tempfile; sys.path.insert(0, %r)\n"
        "import numpy as np\n"
        "from pcutil import compare_trees, load_pcb, make_points, oracle_convert, read_tree\n"
        "rng = np.random.default_rng(29)\n"
        "pcb = load_pcb()\n"
        "for cfg, bs, n, ext in (((100, 8, 10.0), 1000, 60000, 70), ((5000, 96, 1000.0), 10000, 300000, 5000)):\n"
        "    pts = make_points(rng.random((n, 3)) * [ext, ext, ext / 10] - [ext / 2, ext / 2, ext / 20])\n"
        "    o = oracle_convert([pts], config=cfg, batch_size=bs)\n"
        "    nodes, mj, st, _ = pcb.convert([pts], config=cfg, batch_size=bs)\n"
        "    assert st.arrivals == o.arrivals and st.hierarchies == o.hierarchies\n"
        "    compare_trees(o.nodes, nodes, 'oracle', 'gpu-waves')\n"
        "    assert mj == o.meta_json and all(o.nodes[k] == nodes[k] for k in nodes)\n"
        "A = make_points(rng.random((30000, 3)) * 60 - 30)\n"
        "B = make_points(rng.random((20000, 3)) * 80 - 40)\n"
        "with tempfile.TemporaryDirectory() as d_o, tempfile.TemporaryDirectory() as d_g:\n"
        "    oracle_convert([A], out_dir=d_o, config=(100, 8, 16.0), batch_size=1000)\n"
        "    oracle_convert([B], out_dir=d_o, batch_size=1000)\n"
        "    pcb.convert([A], config=(100, 8, 16.0), batch_size=1000, out_dir=d_g)\n"
        "    pcb.convert([B], batch_size=1000, out_dir=d_g)\n"
        "    compare_trees(read_tree(d_o), read_tree(d_g), 'oracle-merge', 'gpu-merge-waves')\n"
        "    assert open(d_o + '/metadata.json').read() == open(d_g + '/metadata.json').read()\n"
        "print('waves ok')\n") % os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, PCB_ENABLE_TUNING="1", PCB_WAVE_POINTS="7000")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "waves ok" in r.stdout, r.stdout + r.stderr


def test_level0_sort_path_forced_and_wide_boxes():
    """Level-0 arrivals are made by a single split pass when the level-0 box has at most 512 cells, else by the
    key / radix sort / gather path.  Both must give the oracle's trees: the sort path forced on small boxes
    (PCB_L0_SORT, fresh process), and a cloud whose box is wider than 512 cells in this process."""
    import os
    import subprocess
    import sys
    rng = np.random.default_rng(31)
    wide = make_points(rng.random((60_000, 3)) * [40.0 * 30, 40.0 * 30, 40.0] - [600.0, 600.0, 0.0])
    _check([wide], config=(50, 8, 40.0), bs=1000)   # 30 x 30 x 1 level-0 cells
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import numpy as np\n"
        "from pcutil import compare_trees, load_pcb, make_points, oracle_convert\n"
        "rng = np.random.default_rng(37)\n"
        "for cfg, bs, n, ext in (((100, 8, 10.0), 1000, 40000, 50), ((5000, 96, 1000.0), 10000, 200000, 2500)):\n"
        "    pts = make_points(rng.random((n, 3)) * ext - ext / 2)\n"
        "    o = oracle_convert([pts], config=cfg, batch_size=bs)\n"
        "    nodes, mj, st, _ = load_pcb().convert([pts], config=cfg, batch_size=bs)\n"
        "    compare_trees(o.nodes, nodes, 'oracle', 'gpu-l0sort')\n"
        "    assert mj == o.meta_json and all(o.nodes[k] == nodes[k] for k in nodes)\n"
        "print('l0 sort ok')\n") % os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, PCB_ENABLE_TUNING="1", PCB_L0_SORT="1")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "l0 sort ok" in r.stdout, r.stdout + r.stderr


def test_gui_converter_preset_batch_50000_count_by_delta(tmp_path):
    """Second caller of the boundary: the in-app converter feeds batches of 50 000 points
    (src/plugins/converter.rs:600-606) and advances number_of_points by the per-cell change of the header total
    (:565-573) instead of by batch length.  Fresh run and merge run, limit edges included (a list that is flushed
    reports a NEGATIVE delta in its cell)."""
    pcb = load_pcb()
    rng = np.random.default_rng(53)
    A = make_points(rng.random((180_000, 3)) * [1500, 900, 40] - [700, 400, 5])
    B = make_points(rng.random((120_001, 3)) * [1500, 900, 40] - [700, 400, 5])
    for cfg in ((5000, 96, 1000.0), (64, 12, 400.0)):
        d_o, d_g = str(tmp_path / f"o{cfg[0]}"), str(tmp_path / f"g{cfg[0]}")
        o1 = oracle_convert([A], out_dir=d_o, config=cfg, batch_size=pcb.GUI_BATCH_SIZE)
        _, mj1, st1, _ = pcb.convert([A], d_g, config=cfg, batch_size=pcb.GUI_BATCH_SIZE, flags=pcb.FLAG_COUNT_BY_DELTA)
        assert o1.points_by_delta == len(A) == st1.number_of_points
        compare_trees(read_tree(d_o), read_tree(d_g), "oracle-50k", "gpu-50k")
        assert mj1 == o1.meta_json
        o2 = oracle_convert([B], out_dir=d_o, batch_size=pcb.GUI_BATCH_SIZE)
        _, mj2, st2, _ = pcb.convert([B], d_g, batch_size=pcb.GUI_BATCH_SIZE, flags=pcb.FLAG_COUNT_BY_DELTA)
        assert o2.points_by_delta == len(B) and st2.number_of_points == len(A) + o2.points_by_delta
        compare_trees(read_tree(d_o), read_tree(d_g), "oracle-50k-merge", "gpu-50k-merge")
        assert mj2 == o2.meta_json


def test_trim_caches_between_conversions():
    """pcb_trim_host_cache (pinned buffers, arenas, stream-ordered pool memory) between conversions of a warm process:
    the next conversion re-creates what it needs and gives the same tree."""
    import torch
    pcb = load_pcb()
    lib = pcb.load_library()
    rng = np.random.default_rng(23)
    pts = make_points(rng.random((120_000, 3)) * [300, 200, 40] - [150, 100, 5])
    cfg = (300, 16, 64.0)
    o = oracle_convert([pts], config=cfg, batch_size=4000)
    for round_no in range(3):
        nodes, mj, st, _ = _gpu([pts], config=cfg, bs=4000)
        compare_trees(o.nodes, nodes, "oracle", f"gpu (conversion {round_no})")
        assert mj == o.meta_json
        free0, _ = torch.cuda.mem_get_info()
        lib.pcb_trim_host_cache()
        free1, _ = torch.cuda.mem_get_info()
        assert free1 >= free0, "trimming must not take device memory"
