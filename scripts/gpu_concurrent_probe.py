"""Probe: do two independent builds overlap usefully on one GPU?  Two halves of the C2 cloud, built one after the
other vs. from two host threads (own handle, stream and arenas each).  usage: gpu_concurrent_probe.py [points]"""
import sys, time, pathlib, threading
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests")); sys.path.insert(0, str(ROOT / "point-cloud_b200"))
import torch, synth
from pcutil import load_pcb
pcb = load_pcb()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
halves = [synth.terrain(n // 2, 2, "cuda", extent=(-2000.0, 0.0, -2000.0, 2000.0)),
          synth.terrain(n // 2, 3, "cuda", extent=(0.0, 2000.0, -2000.0, 2000.0))]
torch.cuda.synchronize()


def build(pts, out):
    cv = pcb.Converter(None, device=0, flags=pcb.FLAG_DEVICE_ONLY)
    try:
        cv.begin_file(); cv.add_points_device(pts.data_ptr(), pts.shape[0]); out.append(cv.finish())
    finally:
        cv.close()


for rep in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for h in halves:
        build(h, [])
    torch.cuda.synchronize(); t_seq = time.perf_counter() - t0
    torch.cuda.synchronize(); t0 = time.perf_counter()
    th = [threading.Thread(target=build, args=(h, [])) for h in halves]
    [t.start() for t in th]; [t.join() for t in th]
    torch.cuda.synchronize(); t_con = time.perf_counter() - t0
    print(f"rep {rep}: sequential {t_seq * 1e3:.2f} ms, concurrent {t_con * 1e3:.2f} ms", flush=True)
