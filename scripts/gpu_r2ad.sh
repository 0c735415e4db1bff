#!/bin/bash
# round-2 call AD (--gpus 4): the C2 bench line of the final tree on four GPUs
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus 4 --no-cpu > gpurun_out/r02_bench_n4_final_terrain.json 2> gpurun_out/bench_n4.err; echo "terrain rc=$?"; tail -2 gpurun_out/bench_n4.err
grep "^{" gpurun_out/r02_bench_n4_final_terrain.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['config']['step_ms'], d.get('check'), (d.get('e2e') or {}).get('value'), d['config'].get('ranks',{}).get('ms_per_step'))"
