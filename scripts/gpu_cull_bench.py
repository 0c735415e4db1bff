#!/usr/bin/env python
"""Times pcb_cull_lod (the renderer's cull / LOD compute pass as a CUDA kernel) on the nodes of a converted
terrain cloud: all nodes of one hierarchy go through ONE dispatch (the per-point work depends on the cell's hierarchy
only, the reference dispatches cell by cell).  CUDA events on the launching stream; bytes = 16 B read per point + 16 B
written per kept point.  Prints one JSON line."""
import json
import re
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(ROOT / "point-cloud_b200"))
from pcutil import Node, load_pcb  # noqa: E402
from test_cull import look_at_rh, perspective_rh  # noqa: E402
import synth  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
    pcb = load_pcb()
    dev = "cuda"
    side = (n / 6.25) ** 0.5
    pts = synth.terrain(n, 2, dev, (-side / 2, side / 2, -side / 2, side / 2), surface_seed=2).cpu().numpy()
    nodes, mj, st, _ = pcb.convert([pts])
    nh = st.hierarchies
    hier = np.array([[1000.0 / 2 ** h, 1000.0 / 2 ** h / 96] for h in range(nh)], dtype=np.float32)   # (cell size, spacing)
    per_h = {h: [] for h in range(nh)}
    cells = []
    for path, data in nodes.items():
        m = re.fullmatch(r"h_(\d+)/c_(-?\d+)_(-?\d+)_(-?\d+)\.bin", path)
        c = tuple(int(g) for g in m.groups())
        cells.append(c)
        nd = Node(data)
        per_h[c[0]].append(nd.grid)
        per_h[c[0]] += [v for v in nd.entries.values() if v is not None]
    cells.sort()
    loaded = np.array(cells, dtype=np.int32)
    far = np.array([1e9] + [4000.0 / 2 ** h for h in range(1, nh)], dtype=np.float32)
    wh = (1920, 1080)
    depth = np.ones((wh[1], wh[0]), dtype=np.float32)
    eye = (-0.6 * side, -0.6 * side, 0.35 * side)
    view, proj = look_at_rh(eye, (0.0, 0.0, 0.0), (0, 0, 1)), perspective_rh(1.0, wh[0] / wh[1], 1.0, 4.0 * side)
    d_hier, d_cells = torch.from_numpy(hier).to(dev), torch.from_numpy(loaded).to(dev)
    d_far, d_depth = torch.from_numpy(far).to(dev), torch.from_numpy(depth).to(dev)
    stream = torch.cuda.current_stream().cuda_stream
    bufs = []
    for h in range(nh):
        if per_h[h]:
            rec = torch.from_numpy(np.ascontiguousarray(np.concatenate(per_h[h], axis=0))).to(dev)
            bufs.append((h, rec))
    # ONE output buffer for all dispatches: the counter (instance_count) runs on across them, as in the renderer
    out_all = torch.empty((sum(int(r.shape[0]) for _, r in bufs), 16), dtype=torch.uint8, device=dev)
    d_cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    res = {}
    for sbd in (False, True):
        def once():
            d_cnt.zero_()
            for h, rec in bufs:
                P = pcb.cull_params(view, proj, eye, wh, nh, h, sbd, nh - 1)
                pcb.cull_lod(0, stream, P, d_hier.data_ptr(), d_cells.data_ptr(), len(loaded), d_far.data_ptr(),
                             d_depth.data_ptr(), rec.data_ptr(), rec.shape[0], out_all.data_ptr(), d_cnt.data_ptr())
        for _ in range(3):
            once()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        e0.record()
        for _ in range(reps):
            once()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        total = sum(int(r.shape[0]) for _, r in bufs)
        kept = int(d_cnt.item())
        res["size_by_distance" if sbd else "plain"] = {
            "ms": round(ms, 3), "points": total, "kept": kept, "Mpoints_per_s": round(total / ms / 1e3, 1),
            "GBps": round((16.0 * total + 16.0 * kept) / ms / 1e6, 1)}
    print(json.dumps({"kernel": "k_cull_lod", "cloud": f"C2-shaped terrain, {n} points", "hierarchies": nh,
                      "loaded_cells": len(loaded), "dispatches": len(bufs), "depth": list(wh), **res}))


if __name__ == "__main__":
    main()
