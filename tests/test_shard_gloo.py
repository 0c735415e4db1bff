"""N > 1 host logic on CPU: two gloo ranks route a point stream by level-0 cell (point-cloud_b200/shard.py).
Checks the invariants the per-rank builds rely on; the builds themselves need a GPU (tests/test_gpu_shard.py)."""
import importlib.util
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _load_shard():
    spec = importlib.util.spec_from_file_location("pcb_shard", ROOT / "point-cloud_b200" / "shard.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _cloud(n, seed):
    rng = np.random.default_rng(seed)
    xyz = (rng.random((n, 3)) * [5000, 3000, 200] - [2500, 1500, 50]).astype(np.float32)
    rec = np.zeros((n, 16), dtype=np.uint8)
    rec[:, :12] = xyz.view(np.uint8).reshape(n, 12)
    rec[:, 12:16] = np.arange(n, dtype=np.uint32).view(np.uint8).reshape(n, 4)  # colour = global index
    return rec


def _worker(rank, world, port, n, bs, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shard = _load_shard()
        full = _cloud(n, 5)
        lo, hi = rank * n // world, (rank + 1) * n // world
        pts, bat, table = shard.route(torch.from_numpy(full[lo:hi]), lo, bs, 1000.0)
        total, hmax, blo, bhi = shard.reduce_metadata(pts.shape[0], 3 + rank, [float(rank), 0.0, -1.0],
                                                      [10.0 + rank, 1.0, 2.0])
        torch.save({"pts": pts, "bat": bat, "table": table, "meta": (total, hmax, blo, bhi)}, f"{tmp}/r{rank}.pt")
    finally:
        dist.destroy_process_group()


def test_route_two_ranks(tmp_path):
    n, bs, world = 20_000, 700, 2
    port = 29600 + os.getpid() % 300
    mp.spawn(_worker, args=(world, port, n, bs, str(tmp_path)), nprocs=world, join=True)
    shard = _load_shard()
    full = _cloud(n, 5)
    outs = [torch.load(f"{tmp_path}/r{r}.pt") for r in range(world)]
    keys_full = shard.l0_cell_keys(torch.from_numpy(full), 1000.0)
    seen = []
    for r, o in enumerate(outs):
        pts, bat = o["pts"].numpy(), o["bat"].numpy()
        gidx = pts[:, 12:16].copy().view(np.uint32).reshape(-1)
        assert np.all(np.diff(gidx.astype(np.int64)) > 0), "order of the undivided stream must be kept"
        assert np.array_equal(bat, (gidx // bs).astype(np.int32)), "global batch ordinals"
        assert np.array_equal(pts, full[gidx]), "records travel unchanged"
        cells_here = set(keys_full[torch.from_numpy(gidx.astype(np.int64))].tolist())
        owners = dict(zip(o["table"]["cells"].tolist(), o["table"]["owners"].tolist()))
        assert all(owners[c] == r for c in cells_here), "a level-0 cell lives on exactly one rank"
        seen.append(gidx)
    allidx = np.sort(np.concatenate(seen))
    assert np.array_equal(allidx, np.arange(n)), "every point reaches exactly one rank"
    # the plan balances: no rank above the mean by more than the largest cell
    counts = outs[0]["table"]["counts"]
    assert max(len(s) for s in seen) <= n / world + int(counts.max())
    # identical tables and reduced metadata on both ranks
    assert torch.equal(outs[0]["table"]["owners"], outs[1]["table"]["owners"])
    assert outs[0]["meta"] == outs[1]["meta"] == (n, 4, [0.0, 0.0, -1.0], [11.0, 1.0, 2.0])


def test_plan_is_deterministic_and_emulation_matches():
    shard = _load_shard()
    full = torch.from_numpy(_cloud(9000, 8))
    parts = shard.route_emulated([full[:4000], full[4000:]], 500, 1000.0)
    again = shard.route_emulated([full[:4000], full[4000:]], 500, 1000.0)
    assert all(torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) for a, b in zip(parts, again))
    assert sum(p.shape[0] for p, _ in parts) == 9000
    assert shard.key_to_index(int(shard.l0_cell_keys(full[:1], 1000.0)[0])) == tuple(
        int(np.floor(v / np.float32(1000.0))) for v in full[0, :12].numpy().view(np.float32))


def test_reduce_metadata_single_process_merge_base():
    shard = _load_shard()
    out = shard.reduce_metadata(10, 2, [0, 0, 0], [1, 1, 1], base_points=5, base_bbox=([-1, 0, 0], [0.5, 3, 1]))
    assert out == (15, 2, [-1.0, 0.0, 0.0], [1.0, 3.0, 1.0])
    assert shard.reduce_metadata(0, 1, [0, 0, 0], [0, 0, 0]) == (0, 1, [0.0, 0.0, 0.0], [0.0, 0.0, 0.0])
