"""Host <-> device copy ceiling of the box at N concurrent ranks: what `e2e` can reach at most.
Every rank moves the bench's per-step volume (1.6 GB in, 1.6 GB out) between page-locked host memory and its GPU with
one cudaMemcpyAsync per buffer: H2D alone, D2H alone, both directions at once; all ranks start together (barrier).
usage: python -m torch.distributed.run --nproc-per-node N scripts/gpu_copy_ceiling.py [bytes] > profiles/...json"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 1_600_000_000
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
h_in = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
h_in.fill_(1)
d_in = torch.empty(nbytes, dtype=torch.uint8, device=dev)
d_out = torch.ones(nbytes, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(mode):
    times = []
    for rep in range(5):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if mode in ("h2d", "both"):
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if mode in ("d2h", "both"):
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
    t = sorted(times)[len(times) // 2]
    if world > 1:
        tt = torch.tensor([t], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t = float(tt.item())
    return t


res = {"ranks": world, "bytes_per_rank_per_direction": nbytes, "host_cpus": os.cpu_count()}
for mode in ("h2d", "d2h", "both"):
    t = run(mode)
    moved = nbytes * (2 if mode == "both" else 1)
    res[mode] = {"ms_max_over_ranks": round(t * 1e3, 2), "GBps_per_rank": round(moved / t / 1e9, 1),
                 "GBps_aggregate": round(world * moved / t / 1e9, 1)}
res["e2e_floor_ms"] = {"serial_h2d_then_d2h": round(res["h2d"]["ms_max_over_ranks"] + res["d2h"]["ms_max_over_ranks"], 2),
                       "full_duplex": res["both"]["ms_max_over_ranks"]}
if rank == 0:
    print(json.dumps(res))
if world > 1:
    dist.destroy_process_group()
