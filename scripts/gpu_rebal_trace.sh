#!/bin/bash
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
P=$((500000000 / 8))
PCB_SHARD_TRACE=1 timeout 600 $TR bench.py --gpus $N --no-cpu --e2e-steps 0 --workload blobs --points $P --rebalance 2 --no-check --steps 3 > gpurun_out/rebal_trace_n$N.json 2> gpurun_out/rebal_trace_n$N.err
grep "\[rebalance\]" gpurun_out/rebal_trace_n$N.err | tail -4
grep "^{" gpurun_out/rebal_trace_n$N.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['config'].get('rebalance'))"
