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
