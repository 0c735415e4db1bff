#!/bin/bash
# usage: gpu_multi2.sh N   re-balance runs (C4 forced at level 2, C1-shaped forced deep + auto) and, at N = 8, C5 at 1 B points
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
P=$((500000000 / N)); C=$((100000000 / N))
B="bench.py --gpus $N --no-cpu --e2e-steps 0"
timeout 900 $TR $B --workload blobs --points $P --rebalance 2 > gpurun_out/bench2_n${N}_blobs_rebal2.json 2> gpurun_out/bench2_n${N}_blobs_rebal2.err; echo "blobs-2 rc=$?"
timeout 900 $TR $B --workload cube --points $C --rebalance 12 > gpurun_out/bench2_n${N}_cube_rebal12.json 2> gpurun_out/bench2_n${N}_cube_rebal12.err; echo "cube-12 rc=$?"; grep -m3 "Error\|error" gpurun_out/bench2_n${N}_cube_rebal12.err
timeout 900 $TR $B --workload cube --points $C > gpurun_out/bench2_n${N}_cube_auto.json 2> gpurun_out/bench2_n${N}_cube_auto.err; echo "cube-auto rc=$?"
PCB_SHARD_TRACE=1 timeout 900 $TR $B --workload blobs --points $P --rebalance 2 --no-check --steps 2 > gpurun_out/bench2_n${N}_blobs_rebal2_trace.json 2>/dev/null
PCB_SHARD_TRACE=1 timeout 900 $TR $B --workload cube --points $C --rebalance 12 --no-check --steps 2 > gpurun_out/bench2_n${N}_cube_rebal12_trace.json 2>/dev/null
if [ $N = 8 ]; then
  timeout 900 $TR bench.py --gpus $N --no-cpu --workload city --points 125000000 > gpurun_out/bench2_n${N}_city1b.json 2> gpurun_out/bench2_n${N}_city1b.err; echo "city rc=$?"
fi
python - <<PY
import json
for w in ("blobs_rebal2","cube_rebal12","cube_auto","city1b","blobs_rebal2_trace","cube_rebal12_trace"):
    try:
        txt=open("gpurun_out/bench2_n${N}_%s.json"%w).read(); d=json.loads([l for l in txt.splitlines() if l.startswith("{")][-1])
        print(w, d["value"], d["ms_per_step"], "e2e", d["e2e"] and (d["e2e"]["value"], d["e2e"].get("phases_ms")), d.get("check"), d["config"].get("rebalance"), d["config"].get("ranks",{}).get("arrivals"), d["config"].get("ranks",{}).get("ms_per_step"))
    except Exception as e: print(w,"ERR",e)
PY
