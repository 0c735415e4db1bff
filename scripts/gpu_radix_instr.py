"""Diagnostics: per-phase cycles of k_onesweep (library built with -DPCB_RADIX_INSTR). usage: gpu_radix_instr.py [points]"""
import sys, pathlib, ctypes as C
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests")); sys.path.insert(0, str(ROOT / "point-cloud_b200"))
import torch, synth
from pcutil import load_pcb
pcb = load_pcb()
L = pcb.load_library()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
pts = synth.terrain(n, 2, "cuda")
torch.cuda.synchronize()
names = ["ticket+zero", "load+count", "scan+publish", "ranking", "lookback(t0)", "barrier wait", "scatter"]
for rep in range(2):
    out = (C.c_ulonglong * 16)()
    L.pcb_debug_radix_instr(out)   # reset
    cv = pcb.Converter(None, device=0, flags=pcb.FLAG_DEVICE_ONLY)
    try:
        cv.begin_file(); cv.add_points_device(pts.data_ptr(), pts.shape[0]); st = cv.finish()
    finally:
        cv.close()
    L.pcb_debug_radix_instr(out)
    tiles = out[15] or 1
    tot = sum(out[i] for i in range(7))
    print(f"rep {rep}: tiles {tiles}, cycles/tile {tot / tiles:.0f}, look-back steps/tile {out[14] / tiles:.2f}")
    print("  " + ", ".join(f"{names[i]} {out[i] / tiles:.0f}" for i in range(7)))
