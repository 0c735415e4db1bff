#!/usr/bin/env python
"""Cold `point_converter` process on a 100 M-point PLY on tmpfs, three times, with the library's host timeline
(PCB_TRACE) of the last run: what a CLI user pays, and where."""
import os, subprocess, sys, time, shutil
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "point-cloud_b200")); sys.path.insert(0, str(ROOT / "tests"))
import torch, synth
from bench import write_ply, terrain_extent
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
pts = synth.terrain(n, 2, "cuda", terrain_extent(n), surface_seed=2).cpu().numpy()
torch.cuda.empty_cache()
base = "/dev/shm/pcb_cold"; os.makedirs(base, exist_ok=True)
ply = base + "/cloud.ply"; write_ply(ply, pts); del pts
cli = str(ROOT / "point-cloud_b200" / "point_converter")
for i in range(3):
    env = dict(os.environ)
    if i == 2: env.update(PCB_ENABLE_TUNING="1", PCB_TRACE="1")
    t0 = time.perf_counter()
    r = subprocess.run([cli, "-o", base + "/out", "-f", ply], capture_output=True, text=True, env=env)
    dt = time.perf_counter() - t0
    print(f"run {i}: {dt*1e3:.0f} ms rc={r.returncode}")
    if i == 2:
        lines = [l for l in r.stderr.splitlines() if l.startswith("[pcb]")]
        for l in lines[:6] + ["..."] + lines[-6:]: print("   ", l)
        print("    other stderr:", [l for l in r.stderr.splitlines() if not l.startswith("[pcb]")][:5])
    shutil.rmtree(base + "/out", ignore_errors=True)
shutil.rmtree(base, ignore_errors=True)
