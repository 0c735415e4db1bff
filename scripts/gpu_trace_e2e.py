"""One end-to-end conversion (host buffer -> node images on the host) with the library's host timeline (PCB_TRACE=1).
usage: PCB_TRACE=1 python scripts/gpu_trace_e2e.py [points]"""
import sys, time, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests")); sys.path.insert(0, str(ROOT / "point-cloud_b200"))
import torch, synth
from pcutil import load_pcb
pcb = load_pcb()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
dev_pts = synth.terrain(n, 2, "cuda")
host_t = torch.empty((n, 16), dtype=torch.uint8, pin_memory=True)
host_t.copy_(dev_pts); torch.cuda.synchronize(); del dev_pts
host = host_t.numpy()
for it in range(3):
    print(f"---- conversion {it}", file=sys.stderr, flush=True)
    t0 = time.perf_counter()
    cv = pcb.Converter(None, device=0, flags=pcb.FLAG_NO_FILES)
    cv.begin_file(); cv.add_points(host)
    t1 = time.perf_counter()
    st = cv.finish()
    t2 = time.perf_counter()
    cv.close()
    print(f"add {1e3*(t1-t0):.1f} ms, finish {1e3*(t2-t1):.1f} ms, close {1e3*(time.perf_counter()-t2):.1f} ms; "
          f"h2d {st.h2d_seconds*1e3:.1f} build {st.build_seconds*1e3:.1f} d2h {st.d2h_seconds*1e3:.1f}", file=sys.stderr)
