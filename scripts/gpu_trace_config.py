"""Device build of one synthetic config with the host timeline (PCB_TRACE=1).
usage: PCB_TRACE=1 python scripts/gpu_trace_config.py <cube|blobs|terrain> [points]"""
import sys, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests")); sys.path.insert(0, str(ROOT / "point-cloud_b200"))
import torch, synth
from pcutil import load_pcb
pcb = load_pcb()
kind = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000_000
pts = {"cube": lambda: synth.uniform_cube(n, 1, "cuda"), "blobs": lambda: synth.gaussian_blobs(n, 4, "cuda"),
       "terrain": lambda: synth.terrain(n, 2, "cuda")}[kind]()
torch.cuda.synchronize()
for it in range(2):
    print(f"---- conversion {it}", file=sys.stderr, flush=True)
    cv = pcb.Converter(None, device=0, flags=pcb.FLAG_DEVICE_ONLY)
    try:
        cv.begin_file(); cv.add_points_device(pts.data_ptr(), pts.shape[0]); st = cv.finish()
    finally:
        cv.close()
