"""Multi-GPU path on ONE GPU: G ranks are emulated one after the other (routing by level-0 cell in-process,
then one build per owner with explicit global batch ordinals).  The union of the ranks' node files must be the
single-process result, bit for bit, and the reduced metadata must match."""
import importlib.util
import tempfile

import numpy as np
import pytest
import torch

from pcutil import ROOT, compare_trees, load_pcb, make_points, oracle_convert, read_tree

pytestmark = pytest.mark.gpu


def _shard():
    spec = importlib.util.spec_from_file_location("pcb_shard", ROOT / "point-cloud_b200" / "shard.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _build_sharded(pts, world, cfg, bs, out_dir=None):
    pcb, shard = load_pcb(), _shard()
    dev = torch.device("cuda", 0)
    raw = torch.from_numpy(pts.view(np.uint8).reshape(-1, 16)).to(dev)
    n = raw.shape[0]
    cuts = [r * n // world for r in range(world + 1)]
    parts = shard.route_emulated([raw[cuts[r]:cuts[r + 1]] for r in range(world)], bs, cfg[2])
    nodes, metas = {}, []
    flags = pcb.FLAG_NO_METADATA_FILE if out_dir else 0
    for p, b in parts:
        with pcb.Converter(out_dir, device=0, batch_size=bs, flags=flags, config=cfg) as cv:
            cv.add_points_device_batched(p.data_ptr(), b.data_ptr(), p.shape[0])
            st = cv.finish()
            mine = cv.nodes()
            assert not (set(mine) & set(nodes)), "two ranks produced the same node"
            nodes.update(mine)
            metas.append((p.shape[0], st.hierarchies, list(st.bbox_min), list(st.bbox_max)))
    total, hmax = sum(m[0] for m in metas), max(m[1] for m in metas)
    lo = np.min([m[2] for m in metas if m[0]], axis=0).astype(np.float32)
    hi = np.max([m[3] for m in metas if m[0]], axis=0).astype(np.float32)
    if out_dir:
        pcb.write_metadata(out_dir, total, hmax, lo, hi, cfg)
    return nodes, (total, hmax, lo, hi)


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_build_equals_single(world):
    rng = np.random.default_rng(21)
    pts = make_points(rng.random((120_000, 3)) * [90, 70, 20] - [45, 35, 5])
    cfg, bs = (300, 12, 16.0), 1000
    o = oracle_convert([pts], config=cfg, batch_size=bs)
    nodes, (total, hmax, lo, hi) = _build_sharded(pts, world, cfg, bs)
    compare_trees(o.nodes, nodes, "oracle", f"{world} ranks")
    assert all(o.nodes[k] == nodes[k] for k in nodes)
    assert total == o.number_of_points and hmax == o.hierarchies
    assert np.array_equal(np.concatenate([lo, hi]), o.bbox)


def test_sharded_build_files_and_metadata():
    rng = np.random.default_rng(22)
    pts = make_points(rng.random((60_000, 3)) * [3000, 2500, 40] - [1500, 1200, 0])
    cfg, bs = (5000, 96, 1000.0), 10_000
    with tempfile.TemporaryDirectory() as d_o, tempfile.TemporaryDirectory() as d_g:
        oracle_convert([pts], out_dir=d_o, config=cfg, batch_size=bs)
        _build_sharded(pts, 2, cfg, bs, out_dir=d_g)
        compare_trees(read_tree(d_o), read_tree(d_g), "oracle", "2 ranks")
        assert open(f"{d_o}/metadata.json").read() == open(f"{d_g}/metadata.json").read()


def test_route_device_kernels_match_reference_routing():
    """libpcb's routing kernels (bbox, dense histogram, owner-digit onesweep pack) against the torch routing."""
    pcb, shard = load_pcb(), _shard()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(23)
    pts = make_points(rng.random((300_000, 3)) * [5000, 3000, 200] - [2500, 1500, 50])
    raw = torch.from_numpy(pts.view(np.uint8).reshape(-1, 16)).to(dev)
    n, world, bs, first = raw.shape[0], 4, 777, 123_456
    sp = torch.cuda.current_stream().cuda_stream
    lo, hi, _ = pcb.route_bbox(0, sp, raw.data_ptr(), n, 1000.0)
    xyz = pts.view(np.uint8).reshape(-1, 16)[:, :12].copy().view(np.float32).reshape(-1, 3)
    assert np.array_equal(np.float32(lo), xyz.min(0)) and np.array_equal(np.float32(hi), xyz.max(0))
    origin = [int(np.floor(np.float32(v) / np.float32(1000.0))) for v in lo]
    dims = [int(np.floor(np.float32(v) / np.float32(1000.0))) - o + 1 for v, o in zip(hi, origin)]
    vol = dims[0] * dims[1] * dims[2]
    hist = torch.empty(vol, dtype=torch.int32, device=dev)
    pcb.route_histogram(0, sp, raw.data_ptr(), n, 1000.0, origin, dims, hist.data_ptr())
    idx = np.floor(xyz / np.float32(1000.0)).astype(np.int64) - np.array(origin)
    dense = (idx[:, 2] * dims[1] + idx[:, 1]) * dims[0] + idx[:, 0]
    assert np.array_equal(hist.cpu().numpy(), np.bincount(dense, minlength=vol))
    cells = np.nonzero(np.bincount(dense, minlength=vol))[0]
    owners = shard.plan_owners(torch.from_numpy(cells), torch.from_numpy(np.bincount(dense, minlength=vol)[cells]), world)
    tab = torch.zeros(vol, dtype=torch.uint8)
    tab[torch.from_numpy(cells)] = owners.to(torch.uint8)
    tab_dev = tab.to(dev)
    send_pts = torch.empty((n, 16), dtype=torch.uint8, device=dev)
    send_bat = torch.empty((n,), dtype=torch.int32, device=dev)
    counts, _ = pcb.route_pack(0, sp, raw.data_ptr(), n, first, bs, 1000.0, origin, dims, tab_dev.data_ptr(), world,
                               send_pts.data_ptr(), send_bat.data_ptr())
    torch.cuda.synchronize()
    dest = tab.numpy()[dense]
    order = np.argsort(dest, kind="stable")
    assert counts == np.bincount(dest, minlength=world).tolist()
    assert np.array_equal(send_pts.cpu().numpy(), pts.view(np.uint8).reshape(-1, 16)[order])
    assert np.array_equal(send_bat.cpu().numpy(), ((first + order) // bs).astype(np.int32))


def _finish_arrivals(pcb, shard, arr, cfg, bs, stop=None):
    """One owner's build from an arrivals dict; returns (Converter kept open, nodes, stats)."""
    cv = pcb.Converter(None, device=0, batch_size=bs, config=cfg)
    shard.add_arrivals_to(cv, arr)
    if stop is not None:
        cv.set_stop_level(stop)
    st = cv.finish()
    return cv, cv.nodes(), st


@pytest.mark.parametrize("world", [2, 3, 5])
def test_fused_cell_exchange_equals_oracle(world):
    """pcb_route_scatter_cells (exchange + level-0 split in ONE kernel) + pcb_add_arrivals: every emulated owner receives
    its level-0 arrivals in (cell, stream order) with global batch ordinals and builds from them directly - the union of
    the owners' trees is the oracle's tree.  Many level-0 cells incl. negative indices, ragged slices."""
    pcb, shard = load_pcb(), _shard()
    rng = np.random.default_rng(200 + world)
    pts = make_points(rng.random((150_001, 3)) * [160, 110, 30] - [85, 40, 10])
    cfg, bs = (300, 12, 16.0), 1000     # 11 x 8 x 3 level-0 cells
    o = oracle_convert([pts], config=cfg, batch_size=bs)
    raw = torch.from_numpy(pts.view(np.uint8).reshape(-1, 16)).cuda()
    n = raw.shape[0]
    cuts = [0] + sorted(rng.integers(0, n, world - 1).tolist()) + [n]
    arrs = shard.route_cells_emulated(pcb, [raw[cuts[r]:cuts[r + 1]] for r in range(world)], bs, cfg[2])
    assert sum(a["n"] for a in arrs) == n
    assert max(a["n"] for a in arrs) <= 1.25 * n / world      # greedy plan over ~260 cells
    nodes, total, hmax = {}, 0, 0
    for a in arrs:
        cv, mine, st = _finish_arrivals(pcb, shard, a, cfg, bs)
        assert not (set(mine) & set(nodes)), "two ranks produced the same node"
        nodes.update(mine)
        total += a["n"]
        hmax = max(hmax, st.hierarchies)
        assert list(st.bbox_min) + list(st.bbox_max) == [np.float32(v) for v in a["bbox"]]
        cv.close()
    compare_trees(o.nodes, nodes, "oracle", f"fused exchange, {world} ranks")
    assert total == o.number_of_points and hmax == o.hierarchies
    assert np.array_equal(np.array(arrs[0]["bbox"], dtype=np.float32), o.bbox)


@pytest.mark.parametrize("stop", [1, 2, 3])
def test_partial_build_and_resume_equals_oracle(stop):
    """pcb_set_stop_level + pcb_pending_arrivals + pcb_add_arrivals(level l): building the levels < l on one handle and the
    levels >= l on another from the pending arrivals gives the oracle's tree (the split point of the level-l re-balance)."""
    pcb, shard = load_pcb(), _shard()
    rng = np.random.default_rng(300 + stop)
    pts = make_points(rng.random((120_000, 3)) * [70, 50, 20] - [30, 20, 5])
    cfg, bs = (40, 4, 32.0), 1000
    o = oracle_convert([pts], config=cfg, batch_size=bs)
    assert o.hierarchies > 3, "the cloud must be deeper than the deepest cut"
    dev = torch.device("cuda", 0)
    with pcb.Converter(None, device=0, batch_size=bs, config=cfg) as top:
        top.begin_file()
        top.add_points(pts)
        top.set_stop_level(stop)
        st_top = top.finish()
        nodes = dict(top.nodes())
        assert all(int(k.split("/")[0][2:]) < stop for k in nodes)
        pend = shard.pending_to_arrivals(top, dev)
        assert pend["level"] == stop and pend["n"] > 0
        cv, low, st_low = _finish_arrivals(pcb, shard, pend, cfg, bs)
        assert all(int(k.split("/")[0][2:]) >= stop for k in low)
        nodes.update(low)
        cv.close()
    compare_trees(o.nodes, nodes, "oracle", f"partial build, cut at level {stop}")
    assert st_top.number_of_points == o.number_of_points and max(st_top.hierarchies, st_low.hierarchies) == o.hierarchies


@pytest.mark.parametrize("world,stop", [(2, 1), (4, 2), (3, 3)])
def test_level_l_rebalance_equals_oracle(world, stop):
    """SURVEY 8e "Levels above l" on emulated ranks: ONE dominant level-0 cell (a C1-shaped cloud: the level-0 cut gives all
    the work to one rank).  Owners build the levels < l, the level-l arrivals are re-balanced cell by cell across the
    ranks (same plan as the distributed path), every rank finishes its share: the union is the oracle's tree and the
    level-l arrivals are spread evenly."""
    pcb, shard = load_pcb(), _shard()
    rng = np.random.default_rng(400 + world)
    dense = rng.random((140_000, 3)) * [15.0, 15.0, 6.0]                       # inside level-0 cell (0, 0, 0)
    sparse = rng.random((10_000, 3)) * [60, 60, 10] - [30, 30, 2]
    xyz = np.concatenate([dense, sparse])[rng.permutation(150_000)]
    pts = make_points(xyz)
    cfg, bs = (150, 10, 16.0), 1000
    o = oracle_convert([pts], config=cfg, batch_size=bs)
    raw = torch.from_numpy(pts.view(np.uint8).reshape(-1, 16)).cuda()
    n = raw.shape[0]
    dev = torch.device("cuda", 0)
    cuts = [r * n // world for r in range(world + 1)]
    arrs = shard.route_cells_emulated(pcb, [raw[cuts[r]:cuts[r + 1]] for r in range(world)], bs, cfg[2])
    assert max(a["n"] for a in arrs) > 0.85 * n                               # the level-0 cut cannot balance this cloud
    nodes, tops, pend, hmax = {}, [], [], 0
    for a in arrs:
        cv, mine, st = _finish_arrivals(pcb, shard, a, cfg, bs, stop=stop)
        nodes.update(mine)
        hmax = max(hmax, st.hierarchies)
        tops.append(cv)                                                        # pending arrivals live in the handle
        pend.append(shard.pending_to_arrivals(cv, dev))
    new = shard.rebalance_emulated(pcb, pend)
    tot = sum(a["n"] for a in new)
    assert tot == sum(p["n"] for p in pend) > 0
    if 8 ** stop >= 4 * world:                                                 # enough level-l cells to balance
        assert max(a["n"] for a in new) <= 1.35 * tot / world
    for a in new:
        cv, mine, st = _finish_arrivals(pcb, shard, a, cfg, bs)
        assert not (set(mine) & set(nodes)), "two ranks produced the same node"
        nodes.update(mine)
        hmax = max(hmax, st.hierarchies)
        cv.close()
    for cv in tops:
        cv.close()
    compare_trees(o.nodes, nodes, "oracle", f"level-{stop} re-balance, {world} ranks")
    assert hmax == o.hierarchies
