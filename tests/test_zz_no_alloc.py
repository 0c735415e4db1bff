"""pcb_route_bbox_static (no counterpart in the reference): the bounding box
of a device-resident slice with NO allocation inside the library - the kernel's scratch lives in a static device array.
Checked through the C ABI against the existing, allocating entry point pcb_route_bbox, on two calls in a row (the array
keeps the last call's contents: the second call must not see them).  Integer / byte work: the bar is exact equality.
Collected last on purpose."""
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
from pcutil import load_pcb, make_points  # noqa: E402

pytestmark = pytest.mark.gpu


def _static_bbox(pcb, torch, dev_pts, n, stream):
    out = torch.full((6,), 7.0, dtype=torch.float32, device="cuda")    # stale values must be overwritten
    err = torch.full((1,), 99, dtype=torch.int32, device="cuda")
    nl = pcb.route_bbox_static(0, stream, dev_pts.data_ptr() if n else 0, n, 1000.0, out.data_ptr(), err.data_ptr())
    torch.cuda.synchronize()
    return out.cpu().numpy(), int(err.item()), nl


def test_static_bbox_equals_allocating_entry_point_twice():
    import torch
    pcb = load_pcb()
    rng = np.random.default_rng(61)
    stream = torch.cuda.current_stream().cuda_stream
    # two different clouds in a row: the second box is strictly INSIDE the first, so left-over minima / maxima of the
    # first call in the shared scratch array would show
    clouds = [make_points(rng.random((300_000, 3)) * [9000, 7000, 400] - [4500, 3500, 100]),
              make_points(rng.random((70_001, 3)) * [10, 20, 5] + [100, -200, 3])]
    for rep in range(2):
        for pts in clouds:
            t = torch.from_numpy(pts.view(np.uint8).reshape(-1, 16)).cuda()
            lo, hi, _ = pcb.route_bbox(0, stream, t.data_ptr(), len(pts), 1000.0)            # existing entry point
            out, err, nl = _static_bbox(pcb, torch, t, len(pts), stream)
            assert nl == 3 and err == 0
            assert np.array_equal(out, np.array(lo + hi, dtype=np.float32)), (out, lo, hi)   # exact: integer / byte work
            assert np.array_equal(out[:3], [pts["x"].min(), pts["y"].min(), pts["z"].min()])
            assert np.array_equal(out[3:], [pts["x"].max(), pts["y"].max(), pts["z"].max()])
    # error bits arrive on the device instead of as a status: a NaN coordinate (existing entry point: PCB_ERR_RANGE)
    bad = clouds[1].copy()
    bad["y"][5] = np.nan
    t = torch.from_numpy(bad.view(np.uint8).reshape(-1, 16)).cuda()
    out, err, _ = _static_bbox(pcb, torch, t, len(bad), stream)
    assert err & 1
    with pytest.raises(pcb.PcbError):
        pcb.route_bbox(0, stream, t.data_ptr(), len(bad), 1000.0)
    # and the call after the failing one is clean again
    t = torch.from_numpy(clouds[0].view(np.uint8).reshape(-1, 16)).cuda()
    out, err, _ = _static_bbox(pcb, torch, t, len(clouds[0]), stream)
    lo, hi, _ = pcb.route_bbox(0, stream, t.data_ptr(), len(clouds[0]), 1000.0)
    assert err == 0 and np.array_equal(out, np.array(lo + hi, dtype=np.float32))


def test_static_bbox_empty_and_size_limit():
    import torch
    pcb = load_pcb()
    stream = torch.cuda.current_stream().cuda_stream
    out, err, nl = _static_bbox(pcb, torch, None, 0, stream)
    assert nl == 2 and err == 0
    assert np.all(np.isposinf(out[:3])) and np.all(np.isneginf(out[3:]))
    dummy = torch.zeros(16, dtype=torch.uint8, device="cuda")
    o6 = torch.zeros(6, dtype=torch.float32, device="cuda")
    with pytest.raises(pcb.PcbError) as e:   # beyond the stated maximum: rejected, nothing launched
        pcb.route_bbox_static(0, stream, dummy.data_ptr(), 1 << 30, 1000.0, o6.data_ptr())
    assert e.value.status == -5 and "PCB_BBOX_STATIC_MAX_POINTS" in str(e.value)
