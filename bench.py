#!/usr/bin/env python
"""bench.py — octree/LOD build throughput (BASELINE.json metric) on B200.

A step = ONE full conversion (all levels) of the synthetic cloud of the workload.
  value  : Mpoints/s with the input resident in HBM and the node images left in HBM (device build)
  e2e    : Mpoints/s through the C ABI with HOST buffers: pinned host points -> H2D -> build -> node
           images D2H into host memory (no file system), wall clock
  roofline: dominant kernel class, algorithmic bytes / CUDA-event time vs MEASURED_PEAKS.json hbm_gbs
  cpu_baseline: the CPU oracle (sequential port of the reference's algorithm; the reference CLI itself
           is single-threaded and Rust cannot be built here) on a bounded, same-density sample
`--impl reference` times that CPU oracle alone (all the threads the algorithm can use: it is parallel
only across level-0 cells).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(ROOT / "point-cloud_b200"))

METRIC = "octree_build_throughput"
UNIT = "Mpoints/s"
DENSITY = 6.25  # points per m^2 of the C2 terrain workload (1e8 points over 4 km x 4 km)


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_oracle_run(n_points, extent, threads, seed=2):
    """Times the CPU oracle on a terrain tile of the workload's density.  Returns (Mpts/s, seconds)."""
    import synth
    from pcutil import oracle_convert
    pts = synth.terrain(n_points, seed, "cpu", extent).numpy()
    t0 = time.perf_counter()
    oracle_convert([pts], dump=False, threads=threads)
    dt = time.perf_counter() - t0
    return n_points / dt / 1e6, dt


class ClockSampler:
    """SM clock / throttle-reason samples DURING the timed region, read through NVML in a thread of this
    process (an `nvidia-smi -lms` child was measured to stall the CUDA calls of the timed steps for up to
    500 ms at a time; in-process NVML queries do not).  Falls back to a slow nvidia-smi loop."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index):
        self.index = index
        self.samples = []   # (time, sm_mhz, max_mhz, reason bits)
        self.t0 = self.t1 = None
        self.stop_flag = False
        self.thread = None
        self.mode = None

    def _nvml_loop(self):
        import pynvml as nv
        h = nv.nvmlDeviceGetHandleByIndex(self.index)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((time.time(), float(sm), float(mx), int(rs)))
            except Exception:
                pass
            time.sleep(0.05)

    def _smi_loop(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        bits = [0x8, 0x40, 0x20, 0x4]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
                f = [x.strip() for x in out.strip().split(",")]
                rs = sum(b for b, v in zip(bits, f[2:6]) if v.lower().startswith("active"))
                self.samples.append((time.time(), float(f[0]), float(f[1]), rs))
            except Exception:
                pass
            time.sleep(1.0)

    def start(self):
        import threading
        try:
            import pynvml as nv
            nv.nvmlInit()
            target, self.mode = self._nvml_loop, "nvml"
        except Exception:
            target, self.mode = self._smi_loop, "nvidia-smi"
        self.thread = threading.Thread(target=target, daemon=True)
        self.thread.start()
        for _ in range(100):   # NVML start-up is over once the first sample is in
            if self.samples:
                break
            time.sleep(0.05)

    def mark_start(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def stop(self):
        if not self.thread:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["sampler not started"]}
        self.stop_flag = True
        self.thread.join(timeout=12)
        inside = [x for x in self.samples if self.t0 is None or self.t0 <= x[0] <= (self.t1 or x[0])]
        use = inside or self.samples[-3:]
        reasons = set()
        for x in use:
            for bit, nm in self.REASONS.items():
                if x[3] & bit:
                    reasons.add(nm)
        sm = [x[1] for x in use]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max((x[2] for x in use), default=None),
                "samples": len(inside), "source": self.mode, "reasons": sorted(reasons)}


def workload_name(n, kind="terrain"):
    if kind == "blobs":
        return (f"C4 Gaussian blobs, {n} points per GPU, 64 Zipf-weighted blobs (sigma 1-100 m) shared by all ranks, "
                f"default config (5000/96/1000.0), batch 10000")
    if kind == "city":
        return (f"C5 city scan, {n} points per GPU, one 10 km x 10 km city (ground + 20000 boxes, 2 cm noise) shared by all "
                f"ranks, default config (5000/96/1000.0), batch 10000")
    return (f"C2 terrain, {n} points per GPU, 2.5-D heightfield + noise, default config (5000/96/1000.0), "
            f"batch 10000")


def run_reference(args, rank, world):
    """CPU arm: the oracle port on the host cores.  Rank 0 only."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    # 2 x 2 level-0 cells at full density: the only parallelism the algorithm has is across level-0 cells
    threads = min(4, cores)
    side = 2000.0
    n = int(DENSITY * side * side * args.cpu_scale)
    vals = []
    for _ in range(max(1, min(args.steps, 2))):
        v, dt = cpu_oracle_run(n, (-1000.0, -1000.0 + side, -1000.0, -1000.0 + side), threads)
        vals.append((v, dt))
    v = statistics.median([x[0] for x in vals])
    dt = statistics.median([x[1] for x in vals])
    sample = (f"{n} points of the same terrain generator at the workload's density ({DENSITY}/m^2) over 2x2 "
              f"level-0 cells, {threads} threads (one per level-0 cell)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(v, 4), "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(vals), "warmup": 0, "ms_per_step": round(dt * 1e3, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32+u64", "data": "synthetic",
        "config": {"workload": workload_name(args.points), "points_per_gpu": args.points,
                   "sample_points": n, "sample": sample},
        "cpu_baseline": {"value": round(v, 4), "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": round(v, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--points", type=int, default=100_000_000, help="points per GPU (C2 = 1e8)")
    ap.add_argument("--workload", default="terrain", choices=["terrain", "city", "blobs"],
                    help="terrain = C2 (the bench line); city = C5 scan of a fixed 10 km x 10 km city, blobs = C4 (extra runs)")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-scale", type=float, default=1.0, help="scales the bounded CPU sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-profile", action="store_true", help="debug: no per-kernel CUDA events in the timed steps")
    ap.add_argument("--no-clocks", action="store_true", help="debug: do not run the nvidia-smi sampler")
    ap.add_argument("--host-log", action="store_true", help="debug: per-step host times (open, finish, close) on stderr")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import synth
    from pcutil import load_pcb
    pcb = load_pcb()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n = args.points
    # Weak scaling: the undivided stream is world * n points of the same terrain over an extent that grows with
    # the world size (same density); rank r holds the index range [r*n, (r+1)*n) of it, i.e. random points of the
    # WHOLE extent, so (world-1)/world of them must cross NVLink to reach the owner of their level-0 cell.
    side_x = (n / DENSITY) ** 0.5 * world
    side_y = (n / DENSITY) ** 0.5
    if args.workload == "city":   # C5: every rank holds random samples of the WHOLE city
        pts = synth.city(n, 5, dev, sample_seed=50 + rank)
    elif args.workload == "blobs":   # C4: every rank holds random samples of the same 64 blobs
        pts = synth.gaussian_blobs(n, 4, dev, sample_seed=40 + rank)
    else:
        pts = synth.terrain(n, 2 + rank, dev, (-side_x / 2, side_x / 2, -side_y / 2, side_y / 2), surface_seed=2)
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream()
    shard = None
    if world > 1:
        import importlib.util
        spec = importlib.util.spec_from_file_location("pcb_shard", ROOT / "point-cloud_b200" / "shard.py")
        shard = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(shard)
    route_launches = [0]
    exchange = [""]

    def ingest(cv, src):
        """single GPU: the resident stream as is; N GPUs: level-0 exchange (libpcb routing kernels + NCCL
        all-to-all-v), then the routed points with their global batch ordinals."""
        if world == 1:
            cv.begin_file()
            cv.add_points_device(src.data_ptr(), n)
            return None
        rp, rb, info = shard.route_device(pcb, src, rank * n, 10_000, 1000.0, stream=stream)
        route_launches[0] += info["launches"]
        exchange[0] = info.get("exchange", "")
        cv.add_points_device_batched(rp.data_ptr(), rb.data_ptr(), rp.shape[0])
        return rp, rb   # keep alive until finish

    host_log = []

    def step(profile=not args.no_profile):
        flags = pcb.FLAG_DEVICE_ONLY | (pcb.FLAG_PROFILE if profile else 0)
        t0 = time.perf_counter()
        cv = pcb.Converter(None, device=local_rank, flags=flags)
        try:
            cv.set_stream(stream.cuda_stream)
            keep = ingest(cv, pts)
            t1 = time.perf_counter()
            st = cv.finish()
            t2 = time.perf_counter()
            return st, cv.kernel_stats()
        finally:
            cv.close()
            host_log.append((round((t1 - t0) * 1e3, 2), round((t2 - t1) * 1e3, 2),
                             round((time.perf_counter() - t2) * 1e3, 2)))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if not args.no_clocks:
        sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    route_launches[0] = 0
    sampler.mark_start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches, kstats, last = 0, {}, None
    step_evs = []
    ev0.record(stream)
    for _ in range(args.steps):
        last, ks = step()
        e = torch.cuda.Event(enable_timing=True)
        e.record(stream)
        step_evs.append(e)
        launches += last.gpu_launches
        for k in ks:
            d = kstats.setdefault(k["name"], dict(launches=0, units=0, bytes=0, ms=0.0))
            for f in ("launches", "units", "bytes", "ms"):
                d[f] += k[f]
    ev1.record(stream)
    barrier()
    route_timed = route_launches[0]
    sampler.mark_end()
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1)
    step_ms = [round(a.elapsed_time(b), 2) for a, b in zip([ev0] + step_evs[:-1], step_evs)]
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    # per-rank share of the job (points owned after the exchange, arrivals over all levels, own device time)
    ranks_info = None
    if world > 1:
        mine = torch.tensor([float(last.points_added), float(last.arrivals), ev0.elapsed_time(ev1) / args.steps],
                            device=dev, dtype=torch.float64)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        ranks_info = {"points_owned": [int(x[0].item()) for x in allr],
                      "arrivals": [int(x[1].item()) for x in allr],
                      "ms_per_step": [round(float(x[2].item()), 2) for x in allr]}
    if args.host_log:
        print("host_log (open, finish, close) ms:", host_log[-args.steps:], file=sys.stderr)
    value = n * world / (ms_per_step * 1e-3) / 1e6

    # ---- e2e: host buffers through the C ABI (H2D + build + D2H of every node image)
    e2e = None
    if args.e2e_steps > 0:
        host = torch.empty((n, 16), dtype=torch.uint8, pin_memory=True)
        host.copy_(pts)
        torch.cuda.synchronize()
        hnp = host.numpy()
        stage = torch.empty_like(pts) if world > 1 else None
        times, d2h = [], 0

        def e2e_step():
            cv = pcb.Converter(None, device=local_rank, flags=pcb.FLAG_NO_FILES)
            try:
                if world == 1:
                    cv.reserve(n)
                    cv.begin_file()
                    cv.add_points(hnp)
                else:   # host slice -> device, exchange, build
                    cv.set_stream(stream.cuda_stream)
                    stage.copy_(host, non_blocking=True)
                    keep = ingest(cv, stage)
                st = cv.finish()
                return st
            finally:
                cv.close()

        e2e_step()  # warm-up (pinned staging cache, memory pool)
        barrier()
        for _ in range(args.e2e_steps):
            t0 = time.perf_counter()
            st = e2e_step()
            torch.cuda.synchronize()
            times.append(time.perf_counter() - t0)
            d2h = int(st.node_bytes)
        t = statistics.median(times)
        if world > 1:
            tt = torch.tensor([t], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt.item())
        e2e = {"value": round(n * world / t / 1e6, 3), "unit": UNIT, "h2d_bytes_per_step": 16 * n,
               "d2h_bytes_per_step": d2h, "ms_per_step": round(t * 1e3, 2)}
        del host

    if rank == 0:
        peak, peak_src = peaks()
        top = max(kstats.items(), key=lambda kv: kv[1]["ms"]) if kstats else None
        roofline = None
        if top and top[1]["ms"] > 0:
            name, d = top
            achieved = d["bytes"] / (d["ms"] * 1e-3) / 1e9
            traffic = None
            tp = ROOT / "profiles" / "traffic.json"
            if tp.exists():
                traffic = json.loads(tp.read_text()).get(name)
            roofline = {"bound": "hbm", "kernel": name, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                        "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                        "launches": d["launches"], "avg_ms": round(d["ms"] / max(1, d["launches"]), 4),
                        "step_share": round(d["ms"] / ms, 4),
                        "kernels": {k: {"ms_per_step": round(v["ms"] / args.steps, 3),
                                        "GBps": round(v["bytes"] / max(v["ms"], 1e-9) / 1e6, 1)} for k, v in
                                    sorted(kstats.items(), key=lambda kv: -kv[1]["ms"])}}
        cpu = None
        if not args.no_cpu:
            ncpu = int(DENSITY * 3000.0 * 1000.0 * args.cpu_scale)
            v, dt = cpu_oracle_run(ncpu, (0.0, 3000.0, 0.0, 1000.0), 1)
            cpu = {"value": round(v, 4), "unit": UNIT, "cores": 1, "kind": "port", "seconds": round(dt, 2),
                   "sample": f"{ncpu} points: three level-0 cells (3 km x 1 km) of the same terrain generator at the "
                             f"workload's density ({DENSITY}/m^2); sequential oracle = the reference CLI's own "
                             f"concurrency (single-threaded)"}
        out = {"metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": round(ms_per_step, 3), "higher_is_better": True,
               "scaling": "weak", "vs_baseline": None, "dtype": "f32+u64", "data": "synthetic",
               "config": {"workload": workload_name(n, args.workload), "points_per_gpu": n,
                          "l2": "inputs (1.6 GB/step) larger than L2, no explicit flush",
                          "levels": int(last.levels_built), "arrivals_per_point": round(last.arrivals / n, 3),
                          "nodes": int(last.nodes_written), "sharding": ("none (1 GPU)" if world == 1 else
                                       "level-0 cell owners; exchange inside the timed step = " + exchange[0]), "step_ms": step_ms,
                          **({"ranks": ranks_info} if ranks_info else {})},
               "e2e": e2e, "gpu_launches": int(launches) + route_timed, "roofline": roofline, "cpu_baseline": cpu,
               "clocks": clocks}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
