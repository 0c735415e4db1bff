"""Device build of BASELINE.json's other configs at size (robustness + timing; the bench line itself stays C2).
usage: python scripts/gpu_configs.py [points]  -> one line per config on stdout"""
import sys, time, json, pathlib, tempfile, shutil
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests")); sys.path.insert(0, str(ROOT / "point-cloud_b200"))
import torch, synth
from pcutil import load_pcb
pcb = load_pcb()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000


def device_build(pts, reps=3):
    best, st = None, None
    for _ in range(reps):
        cv = pcb.Converter(None, device=0, flags=pcb.FLAG_DEVICE_ONLY)
        try:
            cv.begin_file()
            cv.add_points_device(pts.data_ptr(), pts.shape[0])
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            st = cv.finish()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        finally:
            cv.close()
        best = dt if best is None else min(best, dt)
    return best, st


for name, gen in (("C1-shaped uniform cube", lambda: synth.uniform_cube(n, 1, "cuda")),
                  ("C4-shaped gaussian blobs", lambda: synth.gaussian_blobs(n, 4, "cuda")),
                  ("C2 terrain", lambda: synth.terrain(n, 2, "cuda")),
                  ("C5-shaped city scan (10 km x 10 km)", lambda: synth.city(n, 5, "cuda"))):
    pts = gen()
    torch.cuda.synchronize()
    dt, st = device_build(pts)
    print(json.dumps({"config": name, "points": n, "ms": round(dt * 1e3, 2), "Mpoints_per_s": round(n / dt / 1e6, 1),
                      "levels": int(st.levels_built), "arrivals_per_point": round(st.arrivals / n, 2),
                      "nodes": int(st.nodes_written), "launches": int(st.gpu_launches)}), flush=True)
    del pts

# C3: merge a second cloud into an existing output (node files on tmpfs), pinned host buffers, files written.
# Run twice: the first pass pays the one-time pinned / device allocations of the process.
m = min(n, 100_000_000)
hosts = {}
for seed in (2, 3):
    t = torch.empty((m, 16), dtype=torch.uint8, pin_memory=True)
    t.copy_(synth.terrain(m, seed, "cuda"))
    torch.cuda.synchronize()
    hosts[seed] = t
for rep in ("cold", "warm"):
    d = tempfile.mkdtemp(dir="/dev/shm")
    try:
        for seed in (2, 3):
            t0 = time.perf_counter()
            cv = pcb.Converter(d, device=0)
            try:
                cv.begin_file(); cv.add_points(hosts[seed].numpy()); st = cv.finish()
            finally:
                cv.close()
            dt = time.perf_counter() - t0
            print(json.dumps({"config": "C3 " + ("build seed 2 (files on tmpfs)" if seed == 2 else "merge seed 3 into it") +
                              ", " + rep, "points": m, "ms": round(dt * 1e3, 1), "Mpoints_per_s": round(m / dt / 1e6, 1),
                              "levels": int(st.levels_built), "nodes": int(st.nodes_written),
                              "write_s": round(st.write_seconds, 3), "build_s": round(st.build_seconds, 3),
                              "h2d_s": round(st.h2d_seconds, 3), "d2h_s": round(st.d2h_seconds, 3),
                              "number_of_points": int(st.number_of_points)}), flush=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)
