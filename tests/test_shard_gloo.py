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


def _pending(rank, seed=13):
    """synthetic pending arrivals of one rank: a few cells of a level with distinct index triples per rank"""
    rng = np.random.default_rng(seed + rank)
    ncell = 5 + 3 * rank
    counts = rng.integers(1, 400, ncell)
    if rank == 0:
        counts[2] = 5000                                   # one dominant cell
    n = int(counts.sum())
    pts = torch.from_numpy(rng.integers(0, 256, (n, 16)).astype(np.uint8))
    pts[:, 12] = rank                                      # tag: source rank
    bat = torch.arange(n, dtype=torch.int32) + 1000 * rank
    cs = torch.from_numpy(np.concatenate([[0], np.cumsum(counts)]).astype(np.int32))
    idx = torch.from_numpy(np.stack([np.full(ncell, rank), np.arange(ncell), np.zeros(ncell, dtype=np.int64)], axis=1).astype(np.int32))
    return {"level": 3, "pts": pts, "bat": bat, "n": n, "ncell": ncell, "cell_start": cs, "cell_idx": idx.reshape(-1),
            "count_points": 0}


def _rebalance_worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, str(ROOT / "tests"))
    from pcutil import load_pcb
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shard, pcb = _load_shard(), load_pcb()
        out = shard.rebalance(pcb, _pending(rank))
        torch.save({k: v for k, v in out.items() if k != "keep"}, f"{tmp}/b{rank}.pt")
    finally:
        dist.destroy_process_group()


def test_level_l_rebalance_two_ranks(tmp_path):
    """Host logic of the level-l re-balance over gloo: every cell moves WHOLE and unchanged (records, batch ordinals,
    index triple) to exactly one rank, the CSR of the receiver matches, the plan is the deterministic greedy one and the
    distributed result equals the one-process emulation."""
    sys.path.insert(0, str(ROOT / "tests"))
    from pcutil import load_pcb
    world = 2
    port = 29900 + os.getpid() % 90
    mp.spawn(_rebalance_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    shard, pcb = _load_shard(), load_pcb()
    src = [_pending(r) for r in range(world)]
    outs = [torch.load(f"{tmp_path}/b{r}.pt") for r in range(world)]
    emu = shard.rebalance_emulated(pcb, src)
    cells_seen = {}
    for r, (o, e) in enumerate(zip(outs, emu)):
        assert o["level"] == 3 and o["n"] == e["n"] and o["ncell"] == e["ncell"]
        for k in ("pts", "bat", "cell_start", "cell_idx"):
            assert torch.equal(o[k].cpu(), e[k].cpu()), k
        cs, ci = o["cell_start"].numpy(), o["cell_idx"].view(-1, 3).numpy()
        assert cs[0] == 0 and cs[-1] == o["n"] and np.all(np.diff(cs) > 0)
        for c in range(o["ncell"]):
            s_rank, s_cell = int(ci[c, 0]), int(ci[c, 1])
            a, b = int(src[s_rank]["cell_start"][s_cell]), int(src[s_rank]["cell_start"][s_cell + 1])
            assert torch.equal(o["pts"][cs[c]:cs[c + 1]], src[s_rank]["pts"][a:b]), "a cell's records travel unchanged, in order"
            assert torch.equal(o["bat"][cs[c]:cs[c + 1]], src[s_rank]["bat"][a:b])
            assert (s_rank, s_cell) not in cells_seen
            cells_seen[(s_rank, s_cell)] = r
    assert len(cells_seen) == sum(s["ncell"] for s in src)
    tot = sum(s["n"] for s in src)
    assert max(o["n"] for o in outs) <= max(5000, 0.6 * tot)   # the dominant cell alone bounds the imbalance


def test_cell_plan_slots_are_a_partition():
    """cell_plan: the slots of all ranks tile every owner's receive buffer exactly once, cells contiguous, ranks in order
    inside a cell (= stream order of the undivided input)."""
    sys.path.insert(0, str(ROOT / "tests"))
    from pcutil import load_pcb
    shard, pcb = _load_shard(), load_pcb()
    rng = np.random.default_rng(3)
    world, vol = 4, 37
    H = rng.integers(0, 90, (world, vol)).astype(np.int64)
    H[:, 5] = 0
    owners = pcb.plan_owners(H.sum(0), world)
    assert np.array_equal(owners, pcb.plan_owners(H.sum(0), world))
    load = [int(H.sum(0)[owners == o].sum()) for o in range(world)]
    assert max(load) - min(load) <= int(H.sum(0).max())
    plans = [shard.cell_plan(H, owners, r) for r in range(world)]
    for o in range(world):
        used = np.zeros(int(plans[0][1][o]), dtype=np.int32)
        for c in np.nonzero(owners == o)[0]:
            for r in range(world):
                used[plans[r][0][c]:plans[r][0][c] + H[r, c]] += 1
                if r:
                    assert plans[r][0][c] == plans[r - 1][0][c] + H[r - 1, c]
        assert np.all(used == 1)
        cells, start = plans[o][2], plans[o][3]
        assert np.array_equal(np.diff(start.astype(np.int64)), H.sum(0)[cells]) and np.all(owners[cells] == o)
