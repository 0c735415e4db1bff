"""Per-phase wall time of the multi-GPU ingest (route_device) - run under torchrun with 2+ ranks."""
import os, sys, time, pathlib, importlib.util
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests")); sys.path.insert(0, str(ROOT / "point-cloud_b200"))
import torch, torch.distributed as dist, synth
from pcutil import load_pcb
pcb = load_pcb()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
spec = importlib.util.spec_from_file_location("pcb_shard", ROOT / "point-cloud_b200" / "shard.py")
shard = importlib.util.module_from_spec(spec); spec.loader.exec_module(shard)
n = 100_000_000
sx = (n / 6.25) ** 0.5 * world; sy = (n / 6.25) ** 0.5
pts = synth.terrain(n, 2 + rank, dev, (-sx / 2, sx / 2, -sy / 2, sy / 2), surface_seed=2)
stream = torch.cuda.current_stream()
for it in range(4):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    rp, rb, info = shard.route_device(pcb, pts, rank * n, 10_000, 1000.0, stream=stream, trace=(it == 3 and rank == 0))
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    cv = pcb.Converter(None, device=lr, flags=pcb.FLAG_DEVICE_ONLY)
    cv.set_stream(stream.cuda_stream)
    cv.add_points_device_batched(rp.data_ptr(), rb.data_ptr(), rp.shape[0])
    st = cv.finish(); torch.cuda.synchronize()
    t2 = time.perf_counter()
    cv.close()
    if rank == 0:
        print(f"iter {it}: route {1e3*(t1-t0):.2f} ms, build {1e3*(t2-t1):.2f} ms", flush=True)
dist.destroy_process_group()
