#!/bin/bash
# round-2 call U (--gpus 8): the C2 bench line of the final tree on eight GPUs
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus 8 --no-cpu > gpurun_out/r02_bench_n8_final_terrain.json 2> gpurun_out/bench_n8.err; echo "terrain rc=$?"; tail -2 gpurun_out/bench_n8.err
grep "^{" gpurun_out/r02_bench_n8_final_terrain.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['config']['step_ms'], d.get('check'), (d.get('e2e') or {}).get('value'), d['config'].get('ranks',{}).get('ms_per_step'))"
