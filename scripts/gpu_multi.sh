#!/bin/bash
# usage: gpu_multi.sh N   (under gpurun --gpus N): p2p test, copy ceiling, bench C2 + blobs (+ cube at N = 8) at N ranks
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 python -m pytest tests/test_gpu_p2p.py -m gpu -x -q > gpurun_out/pytest_p2p_n$N.log 2>&1; echo "p2p rc=$?" >> gpurun_out/pytest_p2p_n$N.log; tail -3 gpurun_out/pytest_p2p_n$N.log
timeout 600 $TR scripts/gpu_copy_ceiling.py > gpurun_out/copy_ceiling_n$N.json 2> gpurun_out/copy_ceiling_n$N.err; cat gpurun_out/copy_ceiling_n$N.json
timeout 900 $TR bench.py --gpus $N --no-cpu > gpurun_out/bench_n${N}_terrain.json 2> gpurun_out/bench_n${N}_terrain.err; echo "terrain rc=$?"; tail -3 gpurun_out/bench_n${N}_terrain.err
P=$((500000000 / N))
timeout 900 $TR bench.py --gpus $N --no-cpu --workload blobs --points $P > gpurun_out/bench_n${N}_blobs.json 2> gpurun_out/bench_n${N}_blobs.err; echo "blobs rc=$?"; tail -3 gpurun_out/bench_n${N}_blobs.err
timeout 900 $TR bench.py --gpus $N --no-cpu --workload blobs --points $P --rebalance off --e2e-steps 0 --no-check > gpurun_out/bench_n${N}_blobs_norebal.json 2> gpurun_out/bench_n${N}_blobs_norebal.err; echo "blobs-off rc=$?"
if [ "$2" = "extra" ]; then
  timeout 900 $TR bench.py --gpus $N --no-cpu --workload blobs --points $P --rebalance 2 --e2e-steps 0 > gpurun_out/bench_n${N}_blobs_rebal2.json 2> gpurun_out/bench_n${N}_blobs_rebal2.err; echo "blobs-2 rc=$?"; tail -3 gpurun_out/bench_n${N}_blobs_rebal2.err
  C=$((100000000 / N))
  timeout 900 $TR bench.py --gpus $N --no-cpu --workload cube --points $C --e2e-steps 0 > gpurun_out/bench_n${N}_cube.json 2> gpurun_out/bench_n${N}_cube.err; echo "cube rc=$?"; tail -3 gpurun_out/bench_n${N}_cube.err
  timeout 900 $TR bench.py --gpus $N --no-cpu --workload cube --points $C --e2e-steps 0 --rebalance off --no-check > gpurun_out/bench_n${N}_cube_norebal.json 2> gpurun_out/bench_n${N}_cube_norebal.err; echo "cube-off rc=$?"
  Y=$((1000000000 / 8))
  timeout 900 $TR bench.py --gpus $N --no-cpu --workload city --points $Y > gpurun_out/bench_n${N}_city.json 2> gpurun_out/bench_n${N}_city.err; echo "city rc=$?"; tail -3 gpurun_out/bench_n${N}_city.err
fi
python - <<PY
import json
for w in ("terrain","blobs","blobs_norebal","blobs_rebal2","cube","cube_norebal","city"):
    try:
        txt=open("gpurun_out/bench_n${N}_%s.json"%w).read(); d=json.loads([l for l in txt.splitlines() if l.startswith("{")][-1])
        print(w, d["value"], d["ms_per_step"], "e2e", d["e2e"] and (d["e2e"]["value"], d["e2e"].get("phases_ms")), d.get("check"), d["config"].get("rebalance"), d["config"].get("ranks",{}).get("arrivals"), d["config"].get("ranks",{}).get("ms_per_step"))
    except Exception as e: print(w,"ERR",e)
PY
