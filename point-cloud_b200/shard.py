"""Multi-GPU sharding of the octree/LOD build: one process per GPU.

The forest decomposes by level-0 cell: a cell's state depends only on the arrivals from its single parent
(converter.rs:124-134) and level-0 cells have none.  So the undivided input stream (rank r holds the index
range [r*n/G, (r+1)*n/G) of it) is routed ONCE, by level-0 cell, to owner ranks; every rank then builds its
subtrees independently and only three scalars are reduced for metadata.json.

  1. level-0 cell of every local point: floor(p / max_cell_size) per axis, IEEE f32 (metadata.rs:100-102)
  2. all-gather of the local (cell, count) tables -> the same global table on every rank
  3. deterministic greedy bin packing (largest cell first, ties by cell id) -> owner rank per cell
  4. ONE all-to-all-v of the 16-byte point records and ONE of their u32 batch ordinals
     (batch = global input index // batch_size, lib.rs:31-35 applied to the undivided stream).
     Chunks are received in source-rank order and every source sends in index order, so each owner holds
     its points in the order of the undivided stream — the arrival order the reference would have seen.
  5. per-rank build (pcb_add_points_device_batched), node files of different ranks are disjoint
  6. all-reduce: number_of_points (sum), hierarchies (max), bounding box (min / max); rank 0 writes
     metadata.json (pcb_write_metadata).

torch / torch.distributed are plumbing here (device memory, NCCL); the backend is whatever process group the
caller initialised (NCCL over NVLink on the GPU box, gloo in the CPU tests of the routing logic).
"""
from __future__ import annotations

import os
import sys

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

BIAS = 1 << 20  # level-0 indices are validated to lie in [-2^20, 2^20) (PCB_ERR_RANGE otherwise)


def l0_cell_keys(pts_u8: torch.Tensor, max_cell_size: float) -> torch.Tensor:
    """int64 key of the level-0 cell of every point; key order == (x, y, z) lexicographic index order."""
    xyz = pts_u8[:, :12].contiguous().view(torch.float32).view(-1, 3)
    s = torch.tensor(max_cell_size, dtype=torch.float32, device=xyz.device)
    idx = torch.floor(xyz / s).to(torch.int64)  # true IEEE division, as metadata.rs:101
    if idx.numel() and (int(idx.min()) < -BIAS or int(idx.max()) >= BIAS):
        raise ValueError("level-0 cell index outside +-2^20")
    b = idx + BIAS
    return (b[:, 0] << 42) | (b[:, 1] << 21) | b[:, 2]


def key_to_index(key: int) -> Tuple[int, int, int]:
    return ((key >> 42) - BIAS, ((key >> 21) & ((1 << 21) - 1)) - BIAS, (key & ((1 << 21) - 1)) - BIAS)


def plan_owners(cells: torch.Tensor, counts: torch.Tensor, world: int) -> torch.Tensor:
    """Greedy bin packing, largest cell first (ties: smaller key first), onto the least loaded rank (ties:
    lower rank).  Pure function of the global table => identical on every rank.  Returns owner per cell, in
    the order of `cells` (sorted ascending)."""
    cells = cells.cpu().tolist()
    counts = counts.cpu().tolist()
    order = sorted(range(len(cells)), key=lambda i: (-counts[i], cells[i]))
    load = [0] * world
    owner = [0] * len(cells)
    for i in order:
        r = min(range(world), key=lambda q: (load[q], q))
        owner[i] = r
        load[r] += counts[i]
    return torch.tensor(owner, dtype=torch.int64)


def global_cell_table(keys: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """All ranks' (cell, count) tables merged: sorted unique cells + total counts (CPU tensors)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    cells, counts = torch.unique(keys, return_counts=True)
    if world == 1:
        return cells.cpu(), counts.cpu()
    dev = keys.device
    m = torch.tensor([cells.numel()], dtype=torch.int64, device=dev)
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, m, group=group)
    cap = max(int(s.item()) for s in sizes)
    buf = torch.zeros((2, max(cap, 1)), dtype=torch.int64, device=dev)
    buf[0, :cells.numel()] = cells
    buf[1, :cells.numel()] = counts
    bufs = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(bufs, buf, group=group)
    allc = torch.cat([b[0, :int(s.item())] for b, s in zip(bufs, sizes)])
    alln = torch.cat([b[1, :int(s.item())] for b, s in zip(bufs, sizes)])
    ucells, inv = torch.unique(allc, return_inverse=True)
    total = torch.zeros(ucells.numel(), dtype=torch.int64, device=dev).scatter_add_(0, inv, alln)
    return ucells.cpu(), total.cpu()


def route(pts_u8: torch.Tensor, first_index: int, batch_size: int, max_cell_size: float, group=None):
    """Steps 1-4.  pts_u8: (n, 16) uint8 slice of the undivided stream starting at global index first_index.
    Returns (points (m,16) uint8, batch ordinals (m,) int32, owner table dict) for this rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    dev = pts_u8.device
    n = pts_u8.shape[0]
    keys = l0_cell_keys(pts_u8, max_cell_size)
    cells, counts = global_cell_table(keys, group)
    owner_of_cell = plan_owners(cells, counts, world)
    batch = ((torch.arange(n, dtype=torch.int64, device=dev) + first_index) // batch_size).to(torch.int32)
    table = {"cells": cells, "counts": counts, "owners": owner_of_cell}
    if world == 1:
        return pts_u8, batch, table
    slot = torch.searchsorted(cells.to(dev), keys)
    dest = owner_of_cell.to(dev)[slot]
    order = torch.sort(dest, stable=True).indices           # keeps index order inside every destination
    send_pts = pts_u8[order].contiguous()
    send_bat = batch[order].contiguous()
    send_counts = torch.bincount(dest, minlength=world).to(torch.int64)
    recv_counts = torch.zeros(world, dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv_counts, send_counts, group=group)
    ins, outs = send_counts.cpu().tolist(), recv_counts.cpu().tolist()
    m = int(sum(outs))
    recv_pts = torch.empty((m, 16), dtype=torch.uint8, device=dev)
    recv_bat = torch.empty((m,), dtype=torch.int32, device=dev)
    dist.all_to_all_single(recv_pts, send_pts, outs, ins, group=group)
    dist.all_to_all_single(recv_bat, send_bat, outs, ins, group=group)
    table["rank"] = rank
    return recv_pts, recv_bat, table


def reduce_metadata(points_added: int, hierarchies: int, bbox_min, bbox_max, base_points: int = 0,
                    base_bbox=None, group=None):
    """Step 6.  base_* = the existing metadata.json (merge mode), identical on every rank.  Ranks that received
    no points do not contribute a box (Aabb::from of nothing is None, converter.rs:97)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    dev = torch.device("cuda", torch.cuda.current_device()) if (
        dist.is_initialized() and dist.get_backend(group) == "nccl") else torch.device("cpu")
    inf = float("inf")
    has = points_added > 0
    lo = torch.tensor(list(bbox_min) if has else [inf] * 3, dtype=torch.float32, device=dev)
    hi = torch.tensor(list(bbox_max) if has else [-inf] * 3, dtype=torch.float32, device=dev)
    cnt = torch.tensor([points_added, hierarchies], dtype=torch.int64, device=dev)
    hmax = cnt[1:].clone()
    total = cnt[:1].clone()
    if world > 1:
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(total, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(hmax, op=dist.ReduceOp.MAX, group=group)
    n_new = int(total.item())
    if base_points and base_bbox is not None:              # converter.rs:100-102 extend_aabb
        blo = torch.tensor(base_bbox[0], dtype=torch.float32, device=dev)
        bhi = torch.tensor(base_bbox[1], dtype=torch.float32, device=dev)
        lo, hi = (torch.minimum(lo, blo), torch.maximum(hi, bhi)) if n_new else (blo, bhi)
    elif n_new == 0:
        lo = torch.zeros(3, dtype=torch.float32)
        hi = torch.zeros(3, dtype=torch.float32)
    return base_points + n_new, int(hmax.item()), lo.cpu().tolist(), hi.cpu().tolist()


def route_emulated(slices: List[torch.Tensor], batch_size: int, max_cell_size: float):
    """The same routing for G ranks emulated in ONE process (tests on a single GPU / CPU): `slices` are the
    consecutive index ranges of the undivided stream.  Returns [(points, batch ordinals)] per owner rank."""
    world = len(slices)
    keys = [l0_cell_keys(s, max_cell_size) for s in slices]
    cells, counts = torch.unique(torch.cat(keys), return_counts=True)
    owners = plan_owners(cells, counts, world)
    first, out_p, out_b = 0, [[] for _ in range(world)], [[] for _ in range(world)]
    for s, k in zip(slices, keys):
        n = s.shape[0]
        batch = ((torch.arange(n, dtype=torch.int64, device=s.device) + first) // batch_size).to(torch.int32)
        dest = owners.to(s.device)[torch.searchsorted(cells.to(s.device), k)]
        for r in range(world):
            sel = dest == r
            out_p[r].append(s[sel])
            out_b[r].append(batch[sel])
        first += n
    return [(torch.cat(p).contiguous(), torch.cat(b).contiguous()) for p, b in zip(out_p, out_b)]


# ---- device path: the partition runs in the library's own kernels (route.cuh) --------------------------------
class _P2PUnavailable(RuntimeError):
    pass


# Fused partition + exchange (pcb_route_scatter): receive buffers live in torch symmetric memory (one allocation per
# rank, peer-mapped into every process of the node), grown on demand and kept for the life of the process.
_P2P = {"enabled": os.environ.get("PCB_ROUTE_P2P", "1") != "0", "why": None, "buf": None, "hdl": None, "cap": 0}


def _p2p_buffers(dev, group, cap):
    """Symmetric receive buffer for `cap` points: [cap x 16 B points][cap x 4 B batch ordinals].  Collective."""
    if _P2P["buf"] is not None and _P2P["cap"] >= cap:
        return _P2P["buf"], _P2P["hdl"], _P2P["cap"]
    # (Re)allocation: `cap` derives from all-gathered counts and the cached capacity has the same history on every rank,
    # so every rank takes this branch together.  The allocation or the rendezvous can still fail on ONE rank only; the
    # ranks therefore agree on the outcome (all-reduce MIN of an ok flag) and fall back to NCCL together or not at all -
    # a rank that fell back alone would enter all_to_all_single while its peers wait in the symmetric-memory barrier.
    why, buf, hdl, ptrs = None, None, None, None
    try:
        import torch.distributed._symmetric_memory as symm
        cap = (int(cap * 1.1) + 4095) // 4096 * 4096
        buf = symm.empty(cap * 20, dtype=torch.uint8, device=dev)
    except Exception as e:   # noqa: BLE001 - anything here means "no peer-mapped memory on this system"
        why = f"{type(e).__name__}: {e}"
    ok = torch.tensor([0 if why else 1], dtype=torch.int32, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
    if not bool(ok.item()):
        raise _P2PUnavailable(why or "symmetric memory unavailable on a peer rank")
    try:   # collective; every rank is here
        hdl = symm.rendezvous(buf, group if group is not None else dist.group.WORLD)
        ptrs = list(hdl.buffer_ptrs)
    except Exception as e:   # noqa: BLE001
        why = f"{type(e).__name__}: {e}"
    ok = torch.tensor([0 if why else 1], dtype=torch.int32, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
    if not bool(ok.item()):
        raise _P2PUnavailable(why or "symmetric-memory rendezvous failed on a peer rank")
    _P2P.update(buf=buf, hdl=hdl, cap=cap, ptrs=ptrs)
    return buf, hdl, cap


def _route_scatter_p2p(pcb, pts_u8, first_index, batch_size, max_cell_size, origin, dims, owner_dev, local_hist, group,
                       st, mark):
    dev = pts_u8.device
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n = pts_u8.shape[0]
    # who sends how much to whom: local per-owner counts from the local histogram, all-gathered (world x world)
    cnt = torch.zeros(world, dtype=torch.int64, device=dev)
    cnt.index_add_(0, owner_dev.to(torch.int64), local_hist.to(torch.int64))
    allc = torch.empty(world * world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(allc, cnt, group=group)
    cmat = allc.view(world, world).cpu()           # cmat[r][o] = points rank r sends to owner o
    recv_tot = cmat.sum(0)
    m = int(recv_tot[rank])
    offs = cmat[:rank].sum(0).tolist() if rank > 0 else [0] * world
    buf, hdl, cap = _p2p_buffers(dev, group, int(recv_tot.max()))
    ptrs = _P2P["ptrs"]
    mark("counts all-gather")
    with torch.cuda.stream(st):
        hdl.barrier(channel=0)                     # every rank is done with the previous contents
        nl = pcb.route_scatter(dev.index, st.cuda_stream, pts_u8.data_ptr(), n, first_index, batch_size, max_cell_size,
                               origin, dims, owner_dev.data_ptr(), world, ptrs, [p + cap * 16 for p in ptrs], offs)
        hdl.barrier(channel=0)                     # every rank's stores have landed
    # NOTE: views of the persistent symmetric buffer - the next route_device call on ANY rank overwrites them (after its
    # entry barrier).  Consume them (pcb_finish) or copy them before routing again.
    recv_pts = buf[:m * 16].view(m, 16)
    recv_bat = buf[cap * 16:cap * 16 + m * 4].view(torch.int32)
    mark("scatter (partition + exchange)")
    return recv_pts, recv_bat, nl, 20 * (n - int(cmat[rank][rank]))


def route_device(pcb, pts_u8: torch.Tensor, first_index: int, batch_size: int, max_cell_size: float, group=None,
                 stream=None, trace=False):
    """Steps 1-4 on the GPU with libpcb's kernels: bounding box -> dense level-0 histogram -> all-reduce ->
    owner plan -> ONE onesweep pass (owner digit) that writes the all-to-all send buffer -> all-to-all-v.
    Returns (points, batch ordinals (int32), info dict with launch count)."""
    import math
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    dev = pts_u8.device
    n = pts_u8.shape[0]
    st = stream if stream is not None else torch.cuda.current_stream(dev)
    sp = st.cuda_stream
    launches = 0
    import time as _time
    _t = [_time.perf_counter()]

    def _mark(name):   # debugging aid: per-phase wall time (synchronises!)
        if trace:
            torch.cuda.synchronize(dev)
            now = _time.perf_counter()
            print(f"[route] {name}: {1e3 * (now - _t[0]):.2f} ms", flush=True)
            _t[0] = now
    lo, hi, nl = pcb.route_bbox(dev.index, sp, pts_u8.data_ptr(), n, max_cell_size)
    launches += nl
    box = torch.tensor(lo + [-v for v in hi], dtype=torch.float32, device=dev)   # min of (lo, -hi)
    if world > 1:
        dist.all_reduce(box, op=dist.ReduceOp.MIN, group=group)
    b = box.cpu().tolist()
    _mark("bbox + all-reduce")
    glo, ghi = b[:3], [-v for v in b[3:]]
    if not all(math.isfinite(v) for v in glo + ghi):       # no points anywhere
        return pts_u8[:0], torch.zeros(0, dtype=torch.int32, device=dev), {"launches": launches, "cells": 0}
    import numpy as np
    s = np.float32(max_cell_size)
    origin = [int(np.floor(np.float32(v) / s)) for v in glo]      # metadata.rs:100-102 at level 0
    dims = [int(np.floor(np.float32(v) / s)) - o + 1 for v, o in zip(ghi, origin)]
    vol = dims[0] * dims[1] * dims[2]
    hist = torch.empty(vol, dtype=torch.int32, device=dev)
    launches += pcb.route_histogram(dev.index, sp, pts_u8.data_ptr(), n, max_cell_size, origin, dims, hist.data_ptr())
    local_hist = hist.clone() if world > 1 else hist
    if world > 1:
        dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
    h = hist.cpu()
    _mark("histogram + all-reduce")
    cells = torch.nonzero(h).flatten()
    owners = plan_owners(cells, h[cells].to(torch.int64), world)
    owner_tab = torch.zeros(vol, dtype=torch.uint8)
    owner_tab[cells] = owners.to(torch.uint8)
    owner_dev = owner_tab.to(dev)
    _mark("owner plan")
    if world > 1 and world <= 8 and _P2P["enabled"]:
        try:
            out = _route_scatter_p2p(pcb, pts_u8, first_index, batch_size, max_cell_size, origin, dims, owner_dev,
                                     local_hist, group, st, _mark)
            launches += out[2]
            return out[0], out[1], {"launches": launches, "cells": int(cells.numel()), "box": (origin, dims),
                                    "exchange": "p2p stores (pcb_route_scatter over symmetric memory)",
                                    "sent_bytes": out[3]}
        except _P2PUnavailable as e:   # no peer access / symmetric memory on this system: NCCL all-to-all below
            _P2P["enabled"] = False
            _P2P["why"] = str(e)
    send_pts = torch.empty((n, 16), dtype=torch.uint8, device=dev)
    send_bat = torch.empty((n,), dtype=torch.int32, device=dev)
    ins, nl = pcb.route_pack(dev.index, sp, pts_u8.data_ptr(), n, first_index, batch_size, max_cell_size, origin, dims,
                             owner_dev.data_ptr(), world, send_pts.data_ptr(), send_bat.data_ptr())
    launches += nl
    _mark("pack")
    info = {"launches": launches, "cells": int(cells.numel()), "box": (origin, dims)}
    if world == 1:
        return send_pts, send_bat, info
    send_counts = torch.tensor(ins, dtype=torch.int64, device=dev)
    recv_counts = torch.zeros(world, dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv_counts, send_counts, group=group)
    outs = recv_counts.cpu().tolist()
    m = int(sum(outs))
    recv_pts = torch.empty((m, 16), dtype=torch.uint8, device=dev)
    recv_bat = torch.empty((m,), dtype=torch.int32, device=dev)
    dist.all_to_all_single(recv_pts, send_pts, outs, ins, group=group)
    dist.all_to_all_single(recv_bat, send_bat, outs, ins, group=group)
    info["sent_bytes"] = 20 * (n - ins[dist.get_rank(group)])
    info["exchange"] = "NCCL all-to-all-v" + (f" (p2p unavailable: {_P2P['why']})" if _P2P.get("why") else "")
    _mark("all-to-all")
    return recv_pts, recv_bat, info


# ---- level-0 exchange fused with the level-0 split (pcb_route_scatter_cells) ------------------------------------------
def _collective_ok(ok: bool, dev, group) -> bool:
    """all ranks agree: True only if every rank says ok (one all-reduce)"""
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    return bool(flag.item())


def cell_plan(H, owners, rank):
    """Pure host logic of the fused exchange (exercised by the gloo tests).  H: (world, vol) per-rank dense level-0
    histograms (numpy int64), owners: (vol,) owner rank per dense cell.  Returns
      dst_base (vol,) uint32: first slot of THIS rank's points of every cell inside its owner's receive buffer,
      recv (world,) int64: points every owner receives,
      my_cells: dense ids of the cells this rank owns (non-empty, ascending), my_start (len + 1,) uint32 their CSR."""
    import numpy as np
    world, vol = H.shape
    total = H.sum(0)
    excl = np.cumsum(H, axis=0) - H                       # points of lower ranks per cell
    start = np.zeros(vol, dtype=np.int64)
    recv = np.zeros(world, dtype=np.int64)
    for o in range(world):
        mine = owners == o
        c = np.cumsum(np.where(mine, total, 0))
        start[mine] = (c - total)[mine]
        recv[o] = int(c[-1]) if vol else 0
    dst_base = (start + excl[rank]).astype(np.uint32)
    sel = np.nonzero((owners == rank) & (total > 0))[0]
    my_start = np.concatenate([start[sel], [recv[rank]]]).astype(np.uint32)
    return dst_base, recv, sel, my_start


def _level0_box(glo, ghi, max_cell_size):
    import numpy as np
    s = np.float32(max_cell_size)
    origin = [int(np.floor(np.float32(v) / s)) for v in glo]      # metadata.rs:100-102 at level 0
    dims = [int(np.floor(np.float32(v) / s)) - o + 1 for v, o in zip(ghi, origin)]
    return origin, dims


def _cell_indices(cells, origin, dims):
    import numpy as np
    idx = np.empty((len(cells), 3), dtype=np.int32)
    idx[:, 0] = origin[0] + cells % dims[0]
    idx[:, 1] = origin[1] + (cells // dims[0]) % dims[1]
    idx[:, 2] = origin[2] + cells // (dims[0] * dims[1])
    return idx


def route_cells(pcb, pts_u8: torch.Tensor, first_index: int, batch_size: int, max_cell_size: float, group=None,
                stream=None):
    """Steps 1-4 as ONE pass over the slice: bounding box -> dense level-0 histogram -> all-gather of the per-rank
    histograms -> owner plan (pcb_plan_owners) -> pcb_route_scatter_cells stores every point, with its global batch
    ordinal + 1, at its final level-0 arrival slot in its owner's receive buffer (symmetric memory, NVLink stores).
    Returns a dict for Converter.add_arrivals(level 0) - or None when the fused path does not apply (level-0 box wider
    than 512 cells, more than 8 ranks, no peer-mapped memory): callers then use route_device.  The decision is
    collective: every rank returns None or none does."""
    import math
    import numpy as np
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    dev = pts_u8.device
    n = pts_u8.shape[0]
    st = stream if stream is not None else torch.cuda.current_stream(dev)
    sp = st.cuda_stream
    if world < 2 or world > 8 or not _P2P["enabled"]:
        return None
    launches = 0
    lo, hi, nl = pcb.route_bbox(dev.index, sp, pts_u8.data_ptr(), n, max_cell_size)
    launches += nl
    box = torch.tensor(lo + [-v for v in hi], dtype=torch.float32, device=dev)   # min of (lo, -hi)
    dist.all_reduce(box, op=dist.ReduceOp.MIN, group=group)
    b = box.cpu().tolist()
    glo, ghi = b[:3], [-v for v in b[3:]]
    if not all(math.isfinite(v) for v in glo + ghi):       # no points anywhere
        return None
    origin, dims = _level0_box(glo, ghi, max_cell_size)
    vol = dims[0] * dims[1] * dims[2]
    if vol > 512:
        return None
    hist = torch.empty(vol, dtype=torch.int32, device=dev)
    launches += pcb.route_histogram(dev.index, sp, pts_u8.data_ptr(), n, max_cell_size, origin, dims, hist.data_ptr())
    allh = torch.empty(world * vol, dtype=torch.int32, device=dev)
    dist.all_gather_into_tensor(allh, hist, group=group)
    H = allh.view(world, vol).cpu().numpy().astype(np.int64)
    owners = pcb.plan_owners(H.sum(0), world)
    dst_base, recv, my_cells, my_start = cell_plan(H, owners, rank)
    try:   # receive buffers in symmetric memory; _p2p_buffers agrees on the outcome collectively
        buf, hdl, cap = _p2p_buffers(dev, group, int(recv.max()))
    except _P2PUnavailable as e:
        _P2P["enabled"] = False
        _P2P["why"] = str(e)
        return None
    ptrs = _P2P["ptrs"]
    tab = torch.from_numpy(np.concatenate([dst_base.view(np.uint8), owners.astype(np.uint8)])).to(dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    m = int(recv[rank])
    cs = torch.from_numpy(my_start.view(np.int32).copy()).to(dev)
    ci = torch.from_numpy(_cell_indices(my_cells, origin, dims).reshape(-1).copy()).to(dev)
    with torch.cuda.stream(st):
        hdl.barrier(channel=0)                     # every rank is done with the previous contents
        launches += pcb.route_scatter_cells(dev.index, sp, pts_u8.data_ptr(), n, first_index, batch_size, max_cell_size,
                                            origin, dims, tab[4 * vol:].data_ptr(), tab[:4 * vol].data_ptr(), world, ptrs,
                                            [p + cap * 16 for p in ptrs], err.data_ptr())
        hdl.barrier(channel=0)                     # every rank's stores have landed
    # NOTE: views of the persistent symmetric buffer - the next routing call on ANY rank overwrites them
    recv_pts = buf[:m * 16].view(m, 16)
    recv_bat = buf[cap * 16:cap * 16 + m * 4].view(torch.int32)
    return {"level": 0, "pts": recv_pts, "bat": recv_bat, "n": m, "cell_start": cs, "cell_idx": ci,
            "ncell": int(len(my_cells)), "bbox": glo + ghi, "count_points": m, "err": err, "launches": launches,
            "plan_recv": recv.tolist(), "plan_max_cell": int(H.sum(0).max()),
            "cells": int((H.sum(0) > 0).sum()), "box": (origin, dims),
            "sent_bytes": 20 * (n - int(H[rank][owners == rank].sum())),
            "exchange": "p2p stores fused with the level-0 split (pcb_route_scatter_cells over symmetric memory)",
            "keep": (tab,)}


def add_arrivals_to(cv, arr):
    """Hands a routing / re-balance result (dict) to a Converter (pcb_add_arrivals)."""
    cv.add_arrivals(arr["level"], arr["pts"].data_ptr() if arr["n"] else 0, arr["bat"].data_ptr() if arr["n"] else 0,
                    arr["n"], arr["cell_start"].data_ptr() if arr["n"] else 0, arr["cell_idx"].data_ptr() if arr["n"] else 0,
                    arr["ncell"], arr.get("bbox"), arr.get("count_points", 0))
    if arr.get("prism_load"):
        cv.set_prism_load(arr["prism_load"])


def route_cells_emulated(pcb, slices: List[torch.Tensor], batch_size: int, max_cell_size: float):
    """route_cells for G ranks emulated in ONE process on one GPU (tests): the same kernels and the same plan; the
    "peer" receive buffers are local tensors.  Returns one arrivals dict per owner rank."""
    import numpy as np
    world = len(slices)
    dev = slices[0].device
    sp = torch.cuda.current_stream(dev).cuda_stream
    los, his = [], []
    for s in slices:
        if s.shape[0]:
            lo, hi, _ = pcb.route_bbox(dev.index, sp, s.data_ptr(), s.shape[0], max_cell_size)
            los.append(lo)
            his.append(hi)
    glo, ghi = np.min(np.array(los, dtype=np.float32), axis=0).tolist(), np.max(np.array(his, dtype=np.float32), axis=0).tolist()
    origin, dims = _level0_box(glo, ghi, max_cell_size)
    vol = dims[0] * dims[1] * dims[2]
    assert vol <= 512
    H = np.zeros((world, vol), dtype=np.int64)
    for r, s in enumerate(slices):
        hist = torch.zeros(vol, dtype=torch.int32, device=dev)
        pcb.route_histogram(dev.index, sp, s.data_ptr(), s.shape[0], max_cell_size, origin, dims, hist.data_ptr())
        H[r] = hist.cpu().numpy()
    owners = pcb.plan_owners(H.sum(0), world)
    plans = [cell_plan(H, owners, r) for r in range(world)]
    recv = plans[0][1]
    bufs_p = [torch.zeros((max(int(recv[o]), 1), 16), dtype=torch.uint8, device=dev) for o in range(world)]
    bufs_b = [torch.zeros((max(int(recv[o]), 1),), dtype=torch.int32, device=dev) for o in range(world)]
    first = 0
    for r, s in enumerate(slices):
        tab = torch.from_numpy(np.concatenate([plans[r][0].view(np.uint8), owners.astype(np.uint8)])).to(dev)
        err = torch.zeros(1, dtype=torch.int32, device=dev)
        pcb.route_scatter_cells(dev.index, sp, s.data_ptr(), s.shape[0], first, batch_size, max_cell_size, origin, dims,
                                tab[4 * vol:].data_ptr(), tab[:4 * vol].data_ptr(), world, [b.data_ptr() for b in bufs_p],
                                [b.data_ptr() for b in bufs_b], err.data_ptr())
        torch.cuda.synchronize(dev)
        assert int(err.item()) == 0
        first += s.shape[0]
    out = []
    for o in range(world):
        _, _, cells, start = plans[o]
        m = int(recv[o])
        out.append({"level": 0, "pts": bufs_p[o][:m], "bat": bufs_b[o][:m], "n": m,
                    "cell_start": torch.from_numpy(start.view(np.int32).copy()).to(dev),
                    "cell_idx": torch.from_numpy(_cell_indices(cells, origin, dims).reshape(-1).copy()).to(dev),
                    "ncell": int(len(cells)), "bbox": glo + ghi, "count_points": m})
    return out


# ---- level-l re-balance (SURVEY 8e "Levels above l"): cells above the cut are built whole on the owner of their level-0
# cell; the arrivals of level l (the pending arrivals of a partial build, pcb_set_stop_level) are then redistributed
# cell by cell - a cell's state depends only on the deliveries of its single parent (converter.rs:124-134), and those
# deliveries travel with their batch ordinals, so every cell is still processed on exactly one GPU with the reference's
# arrival order.
def pending_to_arrivals(cv, dev):
    """The pending arrivals of a partial build as an arrivals dict (views of the handle's device memory: keep the
    Converter open while they are in use)."""
    import ctypes as C
    p = cv.pending()
    n, ncell = int(p.n), int(p.ncell)

    def view(ptr, nbytes, dtype):
        if not nbytes:
            return torch.zeros(0, dtype=dtype, device=dev)
        return _from_ptr(ptr, nbytes, dev).view(dtype)
    return {"level": int(p.level), "n": n, "ncell": ncell,
            "pts": view(p.d_pts, 16 * n, torch.uint8).view(n, 16) if n else torch.zeros((0, 16), dtype=torch.uint8, device=dev),
            "bat": view(p.d_batch_plus1, 4 * n, torch.int32), "cell_start": view(p.d_cell_start, 4 * (ncell + 1) if n else 0, torch.int32),
            "cell_idx": view(p.d_cell_idx, 12 * ncell, torch.int32), "count_points": 0, "prism_load": float(p.prism_load)}


def _from_ptr(ptr, nbytes, dev):
    """uint8 tensor over existing device memory (no copy) through the CUDA array interface"""
    class _Holder:
        pass
    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(h, device=dev)


def _split_for_owners(arr, owners, world):
    """arr: arrivals dict of this rank; owners: new owner per local cell (numpy).  Returns per destination (cell counts,
    local cell ordinals), cells in local order."""
    import numpy as np
    cs = arr["cell_start"].cpu().numpy().astype(np.int64) if arr["n"] else np.zeros(1, dtype=np.int64)
    lens = np.diff(cs)
    out = []
    for o in range(world):
        sel = np.nonzero(owners == o)[0]
        out.append((lens[sel], sel))
    return out


def _pack_cells(pcb, arr, split):
    """The all-to-all send buffer of this rank: the cells for destination 0, then 1, ... back to back (points, batch
    ordinals) and the element count per destination.  On the GPU one pass of pcb_gather_cells; torch indexing on the CPU
    (gloo tests)."""
    import numpy as np
    dev = arr["pts"].device
    sel = np.concatenate([s[1] for s in split]).astype(np.int64) if split else np.zeros(0, dtype=np.int64)
    cnt = np.concatenate([s[0] for s in split]).astype(np.int64) if split else np.zeros(0, dtype=np.int64)
    ins = [int(s[0].sum()) for s in split]
    total = int(cnt.sum())
    if total == 0:
        return arr["pts"][:0], arr["bat"][:0], ins
    out_start = np.concatenate([[0], np.cumsum(cnt)])
    if dev.type == "cuda":
        tab = torch.from_numpy(np.concatenate([sel, out_start]).astype(np.uint32).view(np.int32).copy()).to(dev)
        out_p = torch.empty((total, 16), dtype=torch.uint8, device=dev)
        out_b = torch.empty((total,), dtype=torch.int32, device=dev)
        pcb.gather_cells(dev.index, torch.cuda.current_stream(dev).cuda_stream, arr["pts"].data_ptr(), arr["bat"].data_ptr(),
                         arr["cell_start"].data_ptr(), tab.data_ptr(), tab[len(sel):].data_ptr(), len(sel), total,
                         out_p.data_ptr(), out_b.data_ptr())
        return out_p, out_b, ins
    cs = arr["cell_start"].numpy().astype(np.int64)
    idx = torch.from_numpy(np.concatenate([np.arange(cs[c], cs[c + 1]) for c in sel]) if len(sel) else np.zeros(0, dtype=np.int64))
    return arr["pts"][idx].contiguous(), arr["bat"][idx].contiguous(), ins


def rebalance_emulated(pcb, arrs):
    """Level-l re-balance for G ranks emulated in one process: arrs[r] = pending arrivals of rank r at the same level.
    Returns the new arrivals dict per rank (cells ordered by (source rank, local order))."""
    import numpy as np
    world = len(arrs)
    dev = arrs[0]["pts"].device
    level = max(a["level"] for a in arrs)
    counts = [np.diff(a["cell_start"].cpu().numpy().astype(np.int64)) if a["n"] else np.zeros(0, dtype=np.int64) for a in arrs]
    owners = pcb.plan_owners(np.concatenate(counts) if counts else np.zeros(0), world)
    offs = np.cumsum([0] + [len(c) for c in counts])
    splits = [_split_for_owners(arrs[r], owners[offs[r]:offs[r + 1]], world) for r in range(world)]
    packed = [_pack_cells(pcb, arrs[r], splits[r]) for r in range(world)]
    out = []
    for o in range(world):
        pts, bat = [], []
        for r in range(world):
            sp, sb, ins = packed[r]
            lo = sum(ins[:o])
            pts.append(sp[lo:lo + ins[o]])
            bat.append(sb[lo:lo + ins[o]])
        cnt = np.concatenate([splits[r][o][0] for r in range(world)])
        idx = [arrs[r]["cell_idx"].view(-1, 3)[torch.from_numpy(splits[r][o][1]).to(dev)] for r in range(world)]
        keep = cnt > 0
        cs = np.concatenate([[0], np.cumsum(cnt[keep])]).astype(np.int32)
        ci = torch.cat(idx)[torch.from_numpy(np.nonzero(keep)[0]).to(dev)].contiguous()
        out.append({"level": level, "pts": torch.cat(pts).contiguous(), "bat": torch.cat(bat).contiguous(), "n": int(cnt.sum()),
                    "cell_start": torch.from_numpy(cs).to(dev), "cell_idx": ci.view(-1), "ncell": int(keep.sum()),
                    "count_points": 0, "prism_load": max(float(a.get("prism_load", 0.0)) for a in arrs)})
    return out


def rebalance(pcb, arr, group=None):
    """Level-l re-balance across the ranks of `group` (collective): all-gather of the (cell count) tables -> the same
    greedy plan on every rank (pcb_plan_owners) -> ONE all-to-all-v of the point records and one of their batch ordinals.
    Returns this rank's new arrivals dict."""
    import numpy as np
    import time as _time
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = arr["pts"].device
    trace = os.environ.get("PCB_SHARD_TRACE") is not None and dev.type == "cuda"
    ph, t0 = {}, [_time.perf_counter()]

    def mark(name):
        if trace:
            torch.cuda.synchronize(dev)
            now = _time.perf_counter()
            ph[name] = round((now - t0[0]) * 1e3, 2)
            t0[0] = now
    mine = np.diff(arr["cell_start"].cpu().numpy().astype(np.int64)) if arr["n"] else np.zeros(0, dtype=np.int64)
    ncs = torch.zeros(world, dtype=torch.int64, device=dev)
    nc = torch.tensor([len(mine)], dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(ncs, nc, group=group)
    ncs = ncs.cpu().numpy()
    cap = max(int(ncs.max()), 1)
    pad = torch.zeros(cap, dtype=torch.int64, device=dev)
    pad[:len(mine)] = torch.from_numpy(mine).to(dev)
    allc = torch.zeros(world * cap, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(allc, pad, group=group)
    allc = allc.view(world, cap).cpu().numpy()
    counts = [allc[r, :ncs[r]] for r in range(world)]
    mark("tables")
    owners = pcb.plan_owners(np.concatenate(counts), world)
    offs = np.cumsum([0] + [len(c) for c in counts])
    mark("plan")
    lv = torch.tensor([arr["level"] if arr["n"] else 0], dtype=torch.int64, device=dev)   # ranks without arrivals do not know it
    dist.all_reduce(lv, op=dist.ReduceOp.MAX, group=group)
    hint = torch.tensor([float(arr.get("prism_load", 0.0))], dtype=torch.float32, device=dev)   # record-filter hint
    dist.all_reduce(hint, op=dist.ReduceOp.MAX, group=group)
    arr = dict(arr, level=int(lv.item()))
    split = _split_for_owners(arr, owners[offs[rank]:offs[rank + 1]], world)
    send_pts, send_bat, ins = _pack_cells(pcb, arr, split)
    mark("gather")
    # what every source sends to me: from the global table (no second count exchange)
    outs, cnt_in, idx_send = [], [], []
    for r in range(world):
        sel = np.nonzero(owners[offs[r]:offs[r + 1]] == rank)[0]
        cnt_in.append(counts[r][sel])
        outs.append(int(counts[r][sel].sum()))
    m = int(sum(outs))
    recv_pts = torch.empty((m, 16), dtype=torch.uint8, device=dev)
    recv_bat = torch.empty((m,), dtype=torch.int32, device=dev)
    dist.all_to_all_single(recv_pts, send_pts, outs, ins, group=group)
    dist.all_to_all_single(recv_bat, send_bat, outs, ins, group=group)
    mark("all_to_all")
    # cell indices of the cells I receive: the senders' index triples, exchanged the same way
    ci = arr["cell_idx"].view(-1, 3)
    send_idx = torch.cat([ci[torch.from_numpy(s[1]).to(dev)] for s in split]).contiguous() if len(mine) else ci
    ins_c = [len(s[1]) for s in split]
    outs_c = [len(c) for c in cnt_in]
    recv_idx = torch.empty((int(sum(outs_c)), 3), dtype=torch.int32, device=dev)
    dist.all_to_all_single(recv_idx, send_idx, outs_c, ins_c, group=group)
    cnt = np.concatenate(cnt_in) if cnt_in else np.zeros(0, dtype=np.int64)
    keep = cnt > 0
    cs = np.concatenate([[0], np.cumsum(cnt[keep])]).astype(np.int32)
    mark("csr")
    if trace and rank == 0:
        print(f"[rebalance] {ph}", file=sys.stderr, flush=True)
    return {"level": arr["level"], "pts": recv_pts, "bat": recv_bat, "n": m, "cell_start": torch.from_numpy(cs).to(dev),
            "cell_idx": recv_idx[torch.from_numpy(np.nonzero(keep)[0]).to(dev)].contiguous().view(-1), "ncell": int(keep.sum()),
            "count_points": 0, "moved_bytes": 20 * (sum(ins) - ins[rank]), "prism_load": float(hint.item())}


class ShardedBuild:
    """One multi-GPU conversion step of this rank: level-0 exchange (fused with the level-0 split when it applies, else
    route_device), optional level-l re-balance, build.  Collective: every rank of `group` calls run() together.

    rebalance: "off" | "auto" | int l.  auto: while the arrivals per rank differ by more than 10 % from the
    mean AND the largest cell of the current level still holds more than 1 / (4 G) of the level's arrivals, the owners
    build one more level; then the arrivals of that level are redistributed cell by cell and every rank finishes its
    share.  Balanced clouds (C2, C5) never leave level 0 and pay nothing."""

    def __init__(self, pcb, device_index, batch_size=10_000, max_cell_size=1000.0, group=None, stream=None, flags=0,
                 out_dir=None, rebalance="auto", max_level=4, profile=False, collect_nodes=False):
        self.pcb, self.dev_index, self.bs, self.S0 = pcb, device_index, batch_size, max_cell_size
        self.group, self.stream, self.flags, self.out_dir = group, stream, flags, out_dir
        self.rebalance, self.max_level = rebalance, max_level
        self.collect_nodes = collect_nodes   # keep every handle's node images (tests, bench check) when handles are retired
        self.handles = []
        self._nodes = {}
        self._kstats = {}
        self.info = {}

    def _open(self):
        cv = self.pcb.Converter(self.out_dir, device=self.dev_index, batch_size=self.bs, flags=self.flags)
        if self.stream is not None:
            cv.set_stream(self.stream.cuda_stream)
        self.handles.append(cv)
        return cv

    def _retire(self, cv):
        """A handle whose pending arrivals have been consumed: keep what is asked for, free its arenas."""
        for k in cv.kernel_stats():
            d = self._kstats.setdefault(k["name"], dict(name=k["name"], launches=0, units=0, bytes=0, ms=0.0))
            for f in ("launches", "units", "bytes", "ms"):
                d[f] += k[f]
        if self.collect_nodes:
            self._nodes.update(cv.nodes())
        cv.close()
        self.handles.remove(cv)

    def close(self):
        for cv in list(self.handles):
            cv.close()
        self.handles = []

    def nodes(self):
        out = dict(self._nodes)
        for cv in self.handles:
            out.update(cv.nodes())
        return out

    def _level_stats(self, arr, dev):
        """(max arrivals per rank, mean, largest cell, total) of the current level over all ranks"""
        import numpy as np
        world = dist.get_world_size(self.group)
        if "plan_recv" in arr:   # level 0 after the fused exchange: the plan already knows, no collective
            r = np.array(arr["plan_recv"], dtype=np.float64)
            return r.max(), r.mean(), arr["plan_max_cell"], r.sum()
        big = int(torch.diff(arr["cell_start"]).max().item()) if arr["n"] else 0
        v = torch.tensor([arr["n"], big], dtype=torch.int64, device=dev)
        mx = v.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=self.group)
        sm = v[:1].clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM, group=self.group)
        return float(mx[0].item()), float(sm[0].item()) / world, int(mx[1].item()), float(sm[0].item())

    def run(self, src: torch.Tensor, first_index: int):
        """src: this rank's slice of the undivided stream ((n, 16) uint8 on the device).  Returns a namespace with the
        summed statistics of this rank's handles."""
        from types import SimpleNamespace
        import time as _time
        pcb = self.pcb
        dev = src.device
        trace = os.environ.get("PCB_SHARD_TRACE") is not None   # debugging aid: synchronises between the phases
        phases = {}
        t_mark = [_time.perf_counter()]

        def mark(name):
            if trace:
                torch.cuda.synchronize(dev)
                now = _time.perf_counter()
                phases[name] = phases.get(name, 0.0) + (now - t_mark[0]) * 1e3
                t_mark[0] = now
        world = dist.get_world_size(self.group)
        launches = 0
        arr = route_cells(pcb, src, first_index, self.bs, self.S0, group=self.group, stream=self.stream)
        mark("route")
        stats = []
        if arr is None:   # general path: stream-ordered points + batch ordinals, the owner splits by level-0 cell itself
            rp, rb, info = route_device(pcb, src, first_index, self.bs, self.S0, group=self.group, stream=self.stream)
            launches += info["launches"]
            self.info = {"exchange": info.get("exchange", ""), "rebalance_level": 0}
            cv = self._open()
            cv.add_points_device_batched(rp.data_ptr(), rb.data_ptr(), rp.shape[0])
            self._keep = (rp, rb)
            stats.append(cv.finish())
        else:
            launches += arr["launches"]
            self.info = {"exchange": arr["exchange"], "rebalance_level": 0}
            self._keep = [arr]
            level = 0
            want = self.rebalance
            while True:
                go_deeper = False
                if want == "auto":
                    mx, mean, big, total = self._level_stats(arr, dev)
                    unbalanced = total > 0 and mx > 1.10 * mean
                    go_deeper = unbalanced and big > total / (4.0 * world) and level < self.max_level
                    do_rebalance = unbalanced and not go_deeper and level > 0 and big <= total / world
                elif want == "off":
                    do_rebalance = False
                else:
                    go_deeper = level < int(want)
                    do_rebalance = not go_deeper and level > 0
                if not go_deeper:
                    break
                prev = self.handles[-1] if self.handles else None
                cv = self._open()                       # the owners build exactly one more level
                add_arrivals_to(cv, arr)
                cv.set_stop_level(level + 1)
                stats.append(cv.finish())
                mark("partial_levels")
                if prev is not None:
                    self._retire(prev)                  # its pending arrivals (this level's input) are consumed
                mark("retire")
                arr = pending_to_arrivals(cv, dev)
                arr["level"] = level + 1
                level += 1
            if do_rebalance:
                arr = rebalance(pcb, arr, group=self.group)
                mark("rebalance")
                if self.handles:
                    self._retire(self.handles[-1])      # the re-balanced arrivals are copies
                self._keep.append(arr)
                self.info["rebalance_level"] = level
                self.info["rebalance_moved_bytes"] = arr.get("moved_bytes", 0)
            cv = self._open()
            add_arrivals_to(cv, arr)
            stats.append(cv.finish())
            mark("final_build")
            self.info["build_arrivals_after_cut"] = int(stats[-1].arrivals)
            if trace:
                self.info["phases_ms"] = {k: round(v, 2) for k, v in phases.items()}
        ks = {k: dict(v) for k, v in self._kstats.items()}
        for cv in self.handles:
            for k in cv.kernel_stats():
                d = ks.setdefault(k["name"], dict(name=k["name"], launches=0, units=0, bytes=0, ms=0.0))
                for f in ("launches", "units", "bytes", "ms"):
                    d[f] += k[f]
        tot = SimpleNamespace(
            points_added=sum(int(s.points_added) for s in stats), arrivals=sum(int(s.arrivals) for s in stats),
            gpu_launches=sum(int(s.gpu_launches) for s in stats) + launches,
            levels_built=sum(int(s.levels_built) for s in stats), nodes_written=sum(int(s.nodes_written) for s in stats),
            node_bytes=sum(int(s.node_bytes) for s in stats), hierarchies=max(int(s.hierarchies) for s in stats),
            number_of_points=sum(int(s.points_added) for s in stats),
            build_seconds=sum(s.build_seconds for s in stats), d2h_seconds=sum(s.d2h_seconds for s in stats),
            h2d_seconds=0.0, bbox_min=list(stats[0].bbox_min), bbox_max=list(stats[0].bbox_max))
        return tot, list(ks.values())
