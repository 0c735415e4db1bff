"""Render-time cull / LOD compute pass (reference: src/plugins/render/point/pipelines/compute.wgsl:81-205).
CPU: known answers for the oracle restatement, each derivable by hand from the cited shader lines.  GPU: pcb_cull_lod
against the oracle on the cells of a converted cloud - float work evaluated in the same IEEE order on both sides, so the
tolerance is ZERO (sets of 16-byte records equal); a real wgpu device may legally fuse or reorder and differ in the last
place of a borderline test (parity unpinned: the shader cannot run on this box)."""
import re

import numpy as np
import pytest

from pcutil import Node, load_pcb, make_points, oracle_convert, oracle_cull, sort_records


def look_at_rh(eye, center, up):
    eye, center, up = (np.asarray(v, dtype=np.float32) for v in (eye, center, up))
    f = center - eye
    f = f / np.linalg.norm(f)
    s = np.cross(f, up)
    s = s / np.linalg.norm(s)
    u = np.cross(s, f)
    m = np.eye(4, dtype=np.float32)
    m[0, :3], m[1, :3], m[2, :3] = s, u, -f
    m[0, 3], m[1, 3], m[2, 3] = -s @ eye, -u @ eye, f @ eye
    return m.astype(np.float32)


def perspective_rh(fovy, aspect, near, far):
    """glam Mat4::perspective_rh (depth 0..1, as wgpu wants it)"""
    f = 1.0 / np.tan(fovy / 2.0)
    m = np.zeros((4, 4), dtype=np.float32)
    m[0, 0], m[1, 1] = f / aspect, f
    m[2, 2], m[2, 3] = far / (near - far), near * far / (near - far)
    m[3, 2] = -1.0
    return m


def _params(pcb, eye, center, wh, nh, cell_h, sbd=False, maxh=0, fovy=1.0, near=0.5, far=500.0):
    return pcb.cull_params(look_at_rh(eye, center, (0, 0, 1)), perspective_rh(fovy, wh[0] / wh[1], near, far), eye, wh, nh,
                           cell_h, sbd, maxh)


def test_oracle_known_answers():
    pcb = load_pcb()   # (only the parameter struct: no device call)
    hier = np.array([[100.0, 1.0], [50.0, 0.5], [25.0, 0.25]], dtype=np.float32)
    loaded = np.array([[0, 0, 0, 0], [1, 0, 0, 0], [1, 1, 0, 0], [2, 0, 0, 0]], dtype=np.int32)   # sorted (h, x, y, z)
    far = np.array([1e9, 60.0, 20.0], dtype=np.float32)
    depth1 = np.ones((8, 8), dtype=np.float32)
    eye = (10.0, -40.0, 10.0)
    P = _params(pcb, eye, (10.0, 10.0, 10.0), (8, 8), 3, 0)
    pts = make_points([[10.0, 10.0, 10.0],     # straight ahead: kept; cells (1,0,0,0) and (2,0,0,0) loaded -> hierarchy 2
                       [52.0, 48.0, 10.0],     # 25 degrees to the right (half fov 28.6); (1,1,0,0) loaded, (2,2,1,0) not -> hierarchy 1
                       [10.0, -80.0, 10.0],    # behind the camera: clip.w < 0 -> ndc.z outside [0, 1] (:186)
                       [10.0, 10.0, 300.0],    # far above the view cone: |ndc.y| > 1
                       [30.0, 30.0, 10.0]])    # (1,0,0,0) loaded, (2,1,1,0) not -> hierarchy 1
    out = oracle_cull(P, hier, loaded, far, depth1, pts)
    assert [tuple(np.round([p["x"], p["y"], p["z"]])) for p in out] == [(10, 10, 10), (52, 48, 10), (30, 30, 10)]
    assert [int(p["rgba"][3]) for p in out] == [2, 1, 1]                       # alpha := hierarchy (:201-202)
    assert all(np.array_equal(o["rgba"][:3], pts[i]["rgba"][:3]) for o, i in zip(out, (0, 1, 4)))
    # a cell of the last hierarchy never searches (:102-104)
    P2 = _params(pcb, eye, (10.0, 10.0, 10.0), (8, 8), 3, 2)
    assert [int(p["rgba"][3]) for p in oracle_cull(P2, hier, loaded, far, depth1, pts)] == [2, 2, 2]
    # size by distance (:88-96): distances 50 / 97.5 / 72.8 -> far[2] = 20 never, far[1] = 60 only for the first point,
    # and an override never LOWERS the hierarchy (loop runs while i > own_hierarchy)
    P3 = _params(pcb, eye, (10.0, 10.0, 10.0), (8, 8), 3, 0, sbd=True, maxh=2)
    assert [int(p["rgba"][3]) for p in oracle_cull(P3, hier, loaded[:1], far, depth1, pts)] == [1, 0, 0]
    assert [int(p["rgba"][3]) for p in oracle_cull(P3, hier, loaded, far, depth1, pts)] == [2, 1, 1]
    # depth test (:198): an occluder far in front of everything removes every point; depth 1.0 keeps them
    assert len(oracle_cull(P, hier, loaded, far, np.zeros((8, 8), dtype=np.float32), pts)) == 0
    # ... except points whose own depth is within 3 x their projected radius of the occluder
    view = look_at_rh(eye, (10.0, 10.0, 10.0), (0, 0, 1))
    proj = perspective_rh(1.0, 1.0, 0.5, 500.0)
    clip = proj @ (view @ np.array([10.0, 10.0, 10.0, 1.0], dtype=np.float32))
    z0 = np.float32(clip[2] / clip[3])
    near_occluder = np.full((8, 8), z0 - np.float32(1e-7), dtype=np.float32)
    out = oracle_cull(P, hier, loaded, far, near_occluder, pts[:1])
    assert len(out) == 1


@pytest.mark.gpu
@pytest.mark.parametrize("size_by_distance", [False, True])
def test_gpu_cull_equals_oracle_on_a_converted_cloud(size_by_distance):
    import torch
    pcb = load_pcb()
    rng = np.random.default_rng(71)
    pts = make_points(rng.random((200_000, 3)) * [120, 90, 25] - [50, 40, 5])
    cfg = (400, 16, 64.0)
    o = oracle_convert([pts], config=cfg, batch_size=5000)
    nh = o.hierarchies
    hier = np.array([[cfg[2] / 2 ** h, cfg[2] / 2 ** h / cfg[1]] for h in range(nh)], dtype=np.float32)
    cells = []
    for path in o.nodes:
        m = re.fullmatch(r"h_(\d+)/c_(-?\d+)_(-?\d+)_(-?\d+)\.bin", path)
        cells.append(tuple(int(g) for g in m.groups()))
    cells.sort()
    loaded = np.array([c for k, c in enumerate(cells) if k % 7 != 3], dtype=np.int32)     # not every cell is resident
    assert nh >= 3 and len(loaded) > 20
    far = np.array([1e9] + [120.0 / 2 ** h for h in range(1, nh)], dtype=np.float32)
    wh = (96, 64)
    depth = (0.97 + 0.03 * rng.random((wh[1], wh[0]))).astype(np.float32)
    depth[rng.random((wh[1], wh[0])) < 0.3] = 1.0
    eye = (-70.0, -60.0, 45.0)
    dev = "cuda"
    d_hier, d_cells = torch.from_numpy(hier).to(dev), torch.from_numpy(loaded).to(dev)
    d_far, d_depth = torch.from_numpy(far).to(dev), torch.from_numpy(depth).to(dev)
    stream = torch.cuda.current_stream().cuda_stream
    kept_total, seen_h = 0, set()
    for h, x, y, z in [c for k, c in enumerate(cells) if k % 5 == 0][:40]:
        nd = Node(o.nodes[f"h_{h}/c_{x}_{y}_{z}.bin"])
        rec = np.concatenate([nd.grid] + [v for v in nd.entries.values() if v is not None], axis=0)
        P = _params(pcb, eye, (10.0, 5.0, 5.0), wh, nh, h, sbd=size_by_distance, maxh=nh - 1, far=400.0)
        want = oracle_cull(P, hier, loaded, far, depth, rec)
        d_in = torch.from_numpy(np.ascontiguousarray(rec)).to(dev)
        d_out = torch.zeros_like(d_in)
        d_cnt = torch.zeros(1, dtype=torch.int32, device=dev)
        for rep in range(2):   # the counter accumulates across dispatches, as instance_count does across cells
            assert pcb.cull_lod(0, stream, P, d_hier.data_ptr(), d_cells.data_ptr(), len(loaded), d_far.data_ptr(),
                                d_depth.data_ptr(), d_in.data_ptr(), len(rec), d_out.data_ptr(), d_cnt.data_ptr()) == 1
            torch.cuda.synchronize()
            if rep == 0:
                k = int(d_cnt.item())
                got = d_out[:k].cpu().numpy()
                d_cnt.zero_()
                d_out.zero_()
        assert k == len(want), (h, x, y, z, k, len(want))
        assert np.array_equal(sort_records(got), sort_records(want)), f"cell {(h, x, y, z)}: kept sets differ"
        assert int(d_cnt.item()) == k and np.array_equal(sort_records(d_out[:k].cpu().numpy()), sort_records(want))
        kept_total += k
        seen_h |= set(np.ascontiguousarray(want).view(np.uint8).reshape(-1, 16)[:, 15].tolist())
    assert kept_total > 2000 and len(seen_h) >= 2, "the scene must exercise the frustum, the depth test and several hierarchies"
