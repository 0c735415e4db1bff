"""Fused partition + exchange over peer memory (pcb_route_scatter + torch symmetric memory) against the NCCL
all-to-all-v path: two ranks, two GPUs.  Needs >= 2 devices (skipped on a one-GPU box)."""
import importlib.util
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu


def _load_shard():
    spec = importlib.util.spec_from_file_location("pcb_shard", ROOT / "point-cloud_b200" / "shard.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _cloud(n, seed):
    rng = np.random.default_rng(seed)
    xyz = (rng.random((n, 3)) * [5000, 3000, 200] - [2500, 1500, 50]).astype(np.float32)
    rec = np.zeros((n, 16), dtype=np.uint8)
    rec[:, :12] = xyz.view(np.uint8).reshape(n, 12)
    rec[:, 12:16] = np.arange(n, dtype=np.uint32).view(np.uint8).reshape(n, 4)  # colour = global index
    return rec


def _worker(rank, world, port, n, bs, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, str(ROOT / "tests"))
    from pcutil import load_pcb
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        shard, pcb = _load_shard(), load_pcb()
        full = _cloud(n, 5)
        lo, hi = rank * n // world, (rank + 1) * n // world
        mine = torch.from_numpy(full[lo:hi]).to(dev)
        res = {}
        for rep in range(2):   # twice: the second call reuses the symmetric buffers
            p1, b1, i1 = shard.route_device(pcb, mine, lo, bs, 1000.0)
            torch.cuda.synchronize()
            res["p2p"] = (p1.cpu().clone(), b1.cpu().clone(), i1.get("exchange", ""))
        shard._P2P["enabled"] = False
        p2, b2, i2 = shard.route_device(pcb, mine, lo, bs, 1000.0)
        torch.cuda.synchronize()
        res["nccl"] = (p2.cpu(), b2.cpu(), i2.get("exchange", ""))
        torch.save(res, f"{tmp}/r{rank}.pt")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_scatter_equals_all_to_all(tmp_path):
    n, bs, world = 400_000, 700, 2
    port = 29700 + os.getpid() % 200
    mp.spawn(_worker, args=(world, port, n, bs, str(tmp_path)), nprocs=world, join=True)
    total = 0
    for r in range(world):
        o = torch.load(f"{tmp_path}/r{r}.pt")
        if not o["p2p"][2].startswith("p2p"):
            pytest.skip("symmetric memory unavailable here: " + o["p2p"][2])
        assert o["nccl"][2].startswith("NCCL")
        assert torch.equal(o["p2p"][0], o["nccl"][0]), "points differ between the fused and the NCCL exchange"
        assert torch.equal(o["p2p"][1], o["nccl"][1]), "batch ordinals differ"
        total += o["p2p"][0].shape[0]
    assert total == n
