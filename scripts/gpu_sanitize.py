"""Small conversions through every sort path (plain, segmented, record filter) for compute-sanitizer.
usage: compute-sanitizer --tool memcheck python scripts/gpu_sanitize.py"""
import sys, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests")); sys.path.insert(0, str(ROOT / "point-cloud_b200"))
import numpy as np
from pcutil import compare_trees, load_pcb, make_points, oracle_convert
rng = np.random.default_rng(3)
for cfg, bs, n, ext in (((100, 8, 10.0), 1000, 30000, 40), ((5000, 96, 1000.0), 10000, 120000, 2500)):
    pts = make_points(rng.random((n, 3)) * [ext, ext, ext / 4] - [ext / 2, ext / 2, ext / 8])
    o = oracle_convert([pts], config=cfg, batch_size=bs)
    nodes, mj, st, _ = load_pcb().convert([pts], config=cfg, batch_size=bs)
    compare_trees(o.nodes, nodes, 'oracle', 'gpu')
    assert mj == o.meta_json
print("sanitize run ok")
