"""ctypes binding of libpcb.so (include/pcb.h) — the host-side mirror of the reference's converter API.

Reference interface mirrored (paths relative to /root/reference):
  Converter::new / add_points_batch / Drop   point-converter/src/converter.rs:78-112, 218-246
  convert_from_paths batch loop              point-converter/src/lib.rs:11-60

There is NO CPU fallback: if the CUDA library is missing or no device is present, construction fails.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Dict, Optional

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libpcb.so"

POINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("rgba", "u1", (4,))])

FLAG_NO_FILES = 1
FLAG_DEVICE_ONLY = 2
FLAG_PROFILE = 4
FLAG_NO_METADATA_FILE = 8
FLAG_COUNT_BY_DELTA = 16   # GUI-caller accounting (src/plugins/converter.rs:565-573)
GUI_BATCH_SIZE = 50_000    # src/plugins/converter.rs:600-606

EXPORTS = [
    "pcb_open", "pcb_set_stream", "pcb_begin_file", "pcb_add_points", "pcb_add_points_batch", "pcb_reserve",
    "pcb_add_points_device", "pcb_finish", "pcb_close", "pcb_trim_host_cache", "pcb_last_error", "pcb_node_count", "pcb_node_get",
    "pcb_route_bbox", "pcb_route_bbox_static", "pcb_route_histogram", "pcb_route_pack", "pcb_route_scatter",
    "pcb_add_las_records", "pcb_alloc_host", "pcb_free_host", "pcb_add_arrivals", "pcb_set_stop_level", "pcb_pending_arrivals", "pcb_set_prism_load",
    "pcb_route_scatter_cells", "pcb_plan_owners", "pcb_cull_lod", "pcb_gather_cells",
    "pcb_add_points_device_batched", "pcb_write_metadata", "pcb_get_config", "pcb_metadata_json", "pcb_kernel_stat_count", "pcb_kernel_stat_get", "pcb_version", "pcb_device_count",
]


class PcbOptions(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("batch_size", C.c_uint32),
                ("flags", C.c_uint32), ("cell_point_overflow_limit", C.c_uint32),
                ("sub_grid_dimension", C.c_uint32), ("max_cell_size", C.c_float), ("reserved", C.c_uint32)]


class PcbStats(C.Structure):
    _fields_ = [("number_of_points", C.c_uint64), ("points_added", C.c_uint64), ("arrivals", C.c_uint64),
                ("nodes_written", C.c_uint64), ("node_bytes", C.c_uint64), ("hierarchies", C.c_uint32),
                ("levels_built", C.c_uint32), ("gpu_launches", C.c_uint64), ("h2d_seconds", C.c_double),
                ("build_seconds", C.c_double), ("d2h_seconds", C.c_double), ("write_seconds", C.c_double),
                ("total_seconds", C.c_double), ("bbox_min", C.c_float * 3), ("bbox_max", C.c_float * 3)]


class PcbKernelStat(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("launches", C.c_uint64), ("units", C.c_uint64), ("bytes", C.c_uint64),
                ("ms", C.c_double)]


class PcbNodeInfo(C.Structure):
    _fields_ = [("hierarchy", C.c_uint32), ("index", C.c_int32 * 3), ("offset", C.c_uint64), ("size", C.c_uint64)]


class PcbArrivals(C.Structure):
    _fields_ = [("level", C.c_uint32), ("prism_load", C.c_float), ("n", C.c_uint64), ("ncell", C.c_uint64),
                ("d_pts", C.c_void_p), ("d_batch_plus1", C.c_void_p), ("d_cell_start", C.c_void_p),
                ("d_cell_idx", C.c_void_p)]


class PcbCullParams(C.Structure):
    _fields_ = [("view", C.c_float * 16), ("projection", C.c_float * 16), ("cam_pos", C.c_float * 3),
                ("depth_width", C.c_uint32), ("depth_height", C.c_uint32), ("number_of_hierarchies", C.c_uint32),
                ("size_by_distance", C.c_uint32), ("max_hierarchy", C.c_uint32), ("cell_hierarchy", C.c_uint32)]


class PcbError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"pcb error {status}: {message}")
        self.status = status


_lib = None


def load_library():
    """Loads libpcb.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(f"{LIB_PATH} is missing: build it with __graft_entry__.build() "
                           f"(make -C point-cloud_b200/csrc); there is no CPU fallback")
    L = C.CDLL(str(LIB_PATH))
    vp = C.c_void_p
    L.pcb_open.argtypes = [C.c_char_p, C.POINTER(PcbOptions), C.POINTER(vp)]
    L.pcb_begin_file.argtypes = [vp]
    L.pcb_add_points.argtypes = [vp, vp, C.c_uint64]
    L.pcb_add_points_batch.argtypes = [vp, vp, C.c_uint64]
    L.pcb_add_points_device.argtypes = [vp, vp, C.c_uint64]
    L.pcb_add_points_device_batched.argtypes = [vp, vp, vp, C.c_uint64]
    L.pcb_write_metadata.argtypes = [C.c_char_p, C.c_uint64, C.c_uint32, C.POINTER(C.c_float), C.POINTER(C.c_float),
                                     C.c_uint32, C.c_uint32, C.c_float]
    L.pcb_get_config.argtypes = [vp, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_float)]
    L.pcb_route_bbox.argtypes = [C.c_int, vp, vp, C.c_uint64, C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.pcb_route_bbox_static.argtypes = [C.c_int, vp, vp, C.c_uint64, C.c_float, vp, vp]
    L.pcb_route_histogram.argtypes = [C.c_int, vp, vp, C.c_uint64, C.c_float, C.POINTER(C.c_int32),
                                      C.POINTER(C.c_uint32), vp]
    L.pcb_route_pack.argtypes = [C.c_int, vp, vp, C.c_uint64, C.c_uint64, C.c_uint32, C.c_float, C.POINTER(C.c_int32),
                                 C.POINTER(C.c_uint32), vp, C.c_uint32, vp, vp, C.POINTER(C.c_uint64)]
    L.pcb_route_scatter.argtypes = [C.c_int, vp, vp, C.c_uint64, C.c_uint64, C.c_uint32, C.c_float, C.POINTER(C.c_int32),
                                    C.POINTER(C.c_uint32), vp, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                    C.POINTER(C.c_uint64)]
    L.pcb_reserve.argtypes = [vp, C.c_uint64]
    L.pcb_add_arrivals.argtypes = [vp, C.c_uint32, vp, vp, C.c_uint64, vp, vp, C.c_uint64, C.POINTER(C.c_float), C.c_uint64]
    L.pcb_set_stop_level.argtypes = [vp, C.c_uint32]
    L.pcb_pending_arrivals.argtypes = [vp, C.POINTER(PcbArrivals)]
    L.pcb_set_prism_load.argtypes = [vp, C.c_float]
    L.pcb_route_scatter_cells.argtypes = [C.c_int, vp, vp, C.c_uint64, C.c_uint64, C.c_uint32, C.c_float, C.POINTER(C.c_int32),
                                          C.POINTER(C.c_uint32), vp, vp, C.c_uint32, C.POINTER(C.c_uint64),
                                          C.POINTER(C.c_uint64), vp]
    L.pcb_plan_owners.argtypes = [vp, C.c_uint64, C.c_uint32, vp]
    L.pcb_gather_cells.argtypes = [C.c_int, vp, vp, vp, vp, vp, vp, C.c_uint32, C.c_uint64, vp, vp]
    L.pcb_cull_lod.argtypes = [C.c_int, vp, C.POINTER(PcbCullParams), vp, vp, C.c_uint32, vp, vp, vp, C.c_uint64, vp, vp]
    L.pcb_add_las_records.argtypes = [vp, vp, C.c_uint64, C.c_uint32, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.pcb_alloc_host.argtypes = [C.c_uint64]
    L.pcb_alloc_host.restype = vp
    L.pcb_free_host.argtypes = [vp]
    L.pcb_free_host.restype = None
    L.pcb_set_stream.argtypes = [vp, vp]
    L.pcb_finish.argtypes = [vp, C.POINTER(PcbStats)]
    L.pcb_close.argtypes = [vp]
    L.pcb_close.restype = None
    L.pcb_last_error.argtypes = [vp]
    L.pcb_last_error.restype = C.c_char_p
    L.pcb_node_count.argtypes = [vp]
    L.pcb_node_count.restype = C.c_uint64
    L.pcb_node_get.argtypes = [vp, C.c_uint64, C.POINTER(PcbNodeInfo), C.POINTER(vp)]
    L.pcb_metadata_json.argtypes = [vp, C.c_char_p, C.c_uint64]
    L.pcb_kernel_stat_count.argtypes = [vp]
    L.pcb_kernel_stat_get.argtypes = [vp, C.c_int, C.POINTER(PcbKernelStat)]
    L.pcb_version.restype = C.c_char_p
    _lib = L
    return L


def _as_records(arr) -> np.ndarray:
    a = np.ascontiguousarray(arr)
    if a.dtype != POINT_DTYPE:
        a = a.view(np.uint8).reshape(-1, 16).view(POINT_DTYPE).reshape(-1)
    return a


class Converter:
    """Mirror of the reference's `Converter` (converter.rs:72-246) on the GPU library.

    out_dir=None keeps every node image in host memory (`nodes()`), nothing touches the file system.
    `config=(limit, sub_grid_dimension, max_cell_size)` seeds a FRESH output only, the way a hand-written
    metadata.json would; an existing metadata.json always wins (lib.rs:86-101)."""

    def __init__(self, out_dir: Optional[str] = None, *, device: int = 0, batch_size: int = 0, flags: int = 0,
                 config=None):
        self._L = load_library()
        opt = PcbOptions()
        opt.struct_size = C.sizeof(PcbOptions)
        opt.device = device
        opt.batch_size = batch_size
        opt.flags = flags
        if config is not None:
            opt.cell_point_overflow_limit, opt.sub_grid_dimension, opt.max_cell_size = (
                int(config[0]), int(config[1]), float(config[2]))
        self._h = C.c_void_p()
        rc = self._L.pcb_open(str(out_dir).encode() if out_dir else None, C.byref(opt), C.byref(self._h))
        if rc != 0:
            raise PcbError(rc, (self._L.pcb_last_error(None) or b"").decode())
        self.stats: Optional[PcbStats] = None

    def _check(self, rc):
        if rc != 0:
            raise PcbError(rc, (self._L.pcb_last_error(self._h) or b"").decode())

    def set_stream(self, cuda_stream: int):
        self._check(self._L.pcb_set_stream(self._h, C.c_void_p(cuda_stream)))

    def reserve(self, n: int):
        self._check(self._L.pcb_reserve(self._h, n))

    def begin_file(self):
        self._check(self._L.pcb_begin_file(self._h))

    def add_points(self, arr):
        """Points of the current input file; cut into batches of batch_size like lib.rs:31-35."""
        a = _as_records(arr)
        self._check(self._L.pcb_add_points(self._h, a.ctypes.data if len(a) else None, len(a)))

    def add_points_batch(self, arr):
        """Exactly one batch — Converter::add_points_batch (converter.rs:106-112)."""
        a = _as_records(arr)
        self._check(self._L.pcb_add_points_batch(self._h, a.ctypes.data if len(a) else None, len(a)))

    def add_points_device(self, device_ptr: int, n: int):
        self._check(self._L.pcb_add_points_device(self._h, C.c_void_p(device_ptr), n))

    def add_points_device_batched(self, device_ptr: int, batch_ptr: int, n: int):
        """Routed points with explicit global batch ordinals (multi-GPU ingest)."""
        self._check(self._L.pcb_add_points_device_batched(self._h, C.c_void_p(device_ptr), C.c_void_p(batch_ptr), n))

    def add_arrivals(self, level: int, pts_ptr: int, bat_ptr: int, n: int, cell_start_ptr: int, cell_idx_ptr: int,
                     ncell: int, bbox6=None, count_points: int = 0):
        """Arrivals of one level in (cell, arrival order) with their CSR - pcb_add_arrivals (device pointers)."""
        box = (C.c_float * 6)(*[float(v) for v in bbox6]) if bbox6 is not None else None
        self._check(self._L.pcb_add_arrivals(self._h, level, C.c_void_p(pts_ptr), C.c_void_p(bat_ptr), n,
                                             C.c_void_p(cell_start_ptr), C.c_void_p(cell_idx_ptr), ncell, box, count_points))

    def set_prism_load(self, v: float):
        """Record-filter hint that travels with pending arrivals (PcbArrivals.prism_load)."""
        self._check(self._L.pcb_set_prism_load(self._h, float(v)))

    def set_stop_level(self, level: int):
        """finish() builds the levels < level only; pending() then describes the arrivals of `level`."""
        self._check(self._L.pcb_set_stop_level(self._h, level))

    def pending(self) -> PcbArrivals:
        a = PcbArrivals()
        self._check(self._L.pcb_pending_arrivals(self._h, C.byref(a)))
        return a

    def add_las_records(self, raw, n: int, stride: int, colour_off: int, scale, offset):
        """Raw uncompressed LAS point records, decoded on the device (converter/las.rs:32-41)."""
        raw = np.ascontiguousarray(raw, dtype=np.uint8)
        sc, of = (C.c_double * 3)(*scale), (C.c_double * 3)(*offset)
        self._check(self._L.pcb_add_las_records(self._h, raw.ctypes.data if n else None, n, stride, colour_off, sc, of))

    def config(self):
        a, b, c = C.c_uint32(), C.c_uint32(), C.c_float()
        self._check(self._L.pcb_get_config(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def finish(self) -> PcbStats:
        """Drop for Converter (converter.rs:241-246): build, write node files + metadata.json."""
        st = PcbStats()
        self._check(self._L.pcb_finish(self._h, C.byref(st)))
        self.stats = st
        return st

    def metadata_json(self) -> str:
        n = self._L.pcb_metadata_json(self._h, None, 0)
        buf = C.create_string_buffer(n + 1)
        self._L.pcb_metadata_json(self._h, buf, n + 1)
        return buf.value.decode()

    def nodes(self) -> Dict[str, bytes]:
        """{relative path: node file bytes} of every node this run produced (host copies)."""
        out = {}
        info = PcbNodeInfo()
        img = C.c_void_p()
        for i in range(self._L.pcb_node_count(self._h)):
            self._check(self._L.pcb_node_get(self._h, i, C.byref(info), C.byref(img)))
            path = "h_%d/c_%d_%d_%d.bin" % (info.hierarchy, info.index[0], info.index[1], info.index[2])
            out[path] = C.string_at(img.value, info.size) if img.value else b""
        return out

    def kernel_stats(self):
        out = []
        ks = PcbKernelStat()
        for i in range(self._L.pcb_kernel_stat_count(self._h)):
            self._L.pcb_kernel_stat_get(self._h, i, C.byref(ks))
            out.append(dict(name=ks.name.decode(), launches=ks.launches, units=ks.units, bytes=ks.bytes, ms=ks.ms))
        return out

    def close(self):
        if self._h:
            self._L.pcb_close(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _route_check(rc):
    if rc < 0:
        raise PcbError(rc, (load_library().pcb_last_error(None) or b"").decode())
    return rc


def route_bbox(device, stream, pts_ptr, n, max_cell_size):
    lo, hi = (C.c_float * 3)(), (C.c_float * 3)()
    nl = _route_check(load_library().pcb_route_bbox(device, C.c_void_p(stream), C.c_void_p(pts_ptr), n,
                                                    max_cell_size, lo, hi))
    return list(lo), list(hi), nl


def route_bbox_static(device, stream, pts_ptr, n, max_cell_size, out6_ptr, err_ptr=0):
    """pcb_route_bbox_static: the bounding box without any allocation (scratch in a static device array); outputs stay on
    the device."""
    return _route_check(load_library().pcb_route_bbox_static(device, C.c_void_p(stream), C.c_void_p(pts_ptr), n,
                                                             max_cell_size, C.c_void_p(out6_ptr),
                                                             C.c_void_p(err_ptr) if err_ptr else None))


def route_histogram(device, stream, pts_ptr, n, max_cell_size, origin, dims, hist_ptr):
    o, d = (C.c_int32 * 3)(*origin), (C.c_uint32 * 3)(*dims)
    return _route_check(load_library().pcb_route_histogram(device, C.c_void_p(stream), C.c_void_p(pts_ptr), n,
                                                           max_cell_size, o, d, C.c_void_p(hist_ptr)))


def route_pack(device, stream, pts_ptr, n, first_index, batch_size, max_cell_size, origin, dims, owner_ptr, world,
               send_pts_ptr, send_bat_ptr):
    o, d = (C.c_int32 * 3)(*origin), (C.c_uint32 * 3)(*dims)
    counts = (C.c_uint64 * world)()
    nl = _route_check(load_library().pcb_route_pack(device, C.c_void_p(stream), C.c_void_p(pts_ptr), n, first_index,
                                                    batch_size, max_cell_size, o, d, C.c_void_p(owner_ptr), world,
                                                    C.c_void_p(send_pts_ptr), C.c_void_p(send_bat_ptr), counts))
    return list(counts), nl


def route_scatter(device, stream, pts_ptr, n, first_index, batch_size, max_cell_size, origin, dims, owner_ptr, world,
                  recv_pts_ptrs, recv_bat_ptrs, recv_offsets):
    """pcb_route_scatter: partition + exchange in one kernel over peer-mapped receive buffers (world <= 8)."""
    o, d = (C.c_int32 * 3)(*origin), (C.c_uint32 * 3)(*dims)
    pp = (C.c_uint64 * world)(*[int(v) for v in recv_pts_ptrs])
    pb = (C.c_uint64 * world)(*[int(v) for v in recv_bat_ptrs])
    po = (C.c_uint64 * world)(*[int(v) for v in recv_offsets])
    return _route_check(load_library().pcb_route_scatter(device, C.c_void_p(stream), C.c_void_p(pts_ptr), n, first_index,
                                                         batch_size, max_cell_size, o, d, C.c_void_p(owner_ptr), world,
                                                         pp, pb, po))


def route_scatter_cells(device, stream, pts_ptr, n, first_index, batch_size, max_cell_size, origin, dims, owner_ptr,
                        dst_base_ptr, world, recv_pts_ptrs, recv_bat_ptrs, err_ptr=0):
    """pcb_route_scatter_cells: level-0 exchange fused with the level-0 split (world <= 8, <= 512 level-0 cells)."""
    o, d = (C.c_int32 * 3)(*origin), (C.c_uint32 * 3)(*dims)
    pp = (C.c_uint64 * world)(*[int(v) for v in recv_pts_ptrs])
    pb = (C.c_uint64 * world)(*[int(v) for v in recv_bat_ptrs])
    return _route_check(load_library().pcb_route_scatter_cells(device, C.c_void_p(stream), C.c_void_p(pts_ptr), n, first_index,
                                                               batch_size, max_cell_size, o, d, C.c_void_p(owner_ptr),
                                                               C.c_void_p(dst_base_ptr), world, pp, pb,
                                                               C.c_void_p(err_ptr) if err_ptr else None))


def gather_cells(device, stream, pts_ptr, bat_ptr, cell_start_ptr, sel_ptr, out_start_ptr, nsel, total, out_pts_ptr, out_bat_ptr):
    """pcb_gather_cells: selected cells of one level's arrivals packed back to back (level-l re-balance)."""
    return _route_check(load_library().pcb_gather_cells(device, C.c_void_p(stream), C.c_void_p(pts_ptr), C.c_void_p(bat_ptr),
                                                        C.c_void_p(cell_start_ptr), C.c_void_p(sel_ptr), C.c_void_p(out_start_ptr),
                                                        nsel, total, C.c_void_p(out_pts_ptr), C.c_void_p(out_bat_ptr)))


def plan_owners(counts, world: int) -> np.ndarray:
    """pcb_plan_owners: deterministic greedy owner plan over a histogram -> uint8 owner per cell."""
    c = np.ascontiguousarray(counts, dtype=np.uint64)
    out = np.zeros(len(c), dtype=np.uint8)
    rc = load_library().pcb_plan_owners(c.ctypes.data, len(c), world, out.ctypes.data)
    if rc != 0:
        raise PcbError(rc, "pcb_plan_owners failed")
    return out


def cull_params(view, projection, cam_pos, depth_wh, number_of_hierarchies, cell_hierarchy, size_by_distance=False,
                max_hierarchy=0) -> PcbCullParams:
    """view / projection: 4 x 4 numpy matrices in the usual math layout (M @ v); stored column-major as glam does."""
    p = PcbCullParams()
    p.view = (C.c_float * 16)(*np.asarray(view, dtype=np.float32).T.reshape(-1).tolist())
    p.projection = (C.c_float * 16)(*np.asarray(projection, dtype=np.float32).T.reshape(-1).tolist())
    p.cam_pos = (C.c_float * 3)(*[float(v) for v in cam_pos])
    p.depth_width, p.depth_height = int(depth_wh[0]), int(depth_wh[1])
    p.number_of_hierarchies, p.cell_hierarchy = int(number_of_hierarchies), int(cell_hierarchy)
    p.size_by_distance, p.max_hierarchy = int(bool(size_by_distance)), int(max_hierarchy)
    return p


def cull_lod(device, stream, params, hier_ptr, cells_ptr, n_loaded, far_ptr, depth_ptr, in_ptr, n, out_ptr, count_ptr):
    """pcb_cull_lod: the renderer's per-cell cull / LOD compute pass (compute.wgsl:81-205) on device buffers."""
    return _route_check(load_library().pcb_cull_lod(device, C.c_void_p(stream), C.byref(params), C.c_void_p(hier_ptr),
                                                    C.c_void_p(cells_ptr), n_loaded, C.c_void_p(far_ptr) if far_ptr else None,
                                                    C.c_void_p(depth_ptr), C.c_void_p(in_ptr), n, C.c_void_p(out_ptr),
                                                    C.c_void_p(count_ptr)))


def write_metadata(out_dir, number_of_points, hierarchies, bbox_min, bbox_max, config):
    L = load_library()
    mn, mx = (C.c_float * 3)(*[float(v) for v in bbox_min]), (C.c_float * 3)(*[float(v) for v in bbox_max])
    rc = L.pcb_write_metadata(str(out_dir).encode(), int(number_of_points), int(hierarchies), mn, mx,
                              int(config[0]), int(config[1]), float(config[2]))
    if rc != 0:
        raise PcbError(rc, "pcb_write_metadata failed")


def convert(files, out_dir=None, *, config=None, batch_size=0, device=0, flags=0):
    """convert_from_paths (lib.rs:11-60) over in-memory 'files' (arrays of 16-byte point records).
    Returns (nodes, metadata_json, stats, kernel_stats)."""
    with Converter(out_dir, device=device, batch_size=batch_size, flags=flags, config=config) as cv:
        for f in files:
            cv.begin_file()
            cv.add_points(f)
        st = cv.finish()
        return cv.nodes(), cv.metadata_json(), st, cv.kernel_stats()
