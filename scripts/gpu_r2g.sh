#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shard.py tests/test_gpu_fullsize.py::test_partial_build_at_size_equals_single_build -m gpu -x -q > gpurun_out/pytest_shard.log 2>&1; echo "shard rc=$?" >> gpurun_out/pytest_shard.log
tail -6 gpurun_out/pytest_shard.log
bash scripts/gpu_rebal_trace.sh 2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus 2 --no-cpu --e2e-steps 0 --workload blobs --points 62500000 --rebalance 2 > gpurun_out/bench3_n2_blobs_rebal2.json 2>/dev/null
timeout 600 $TR bench.py --gpus 2 --no-cpu --e2e-steps 0 --workload blobs --points 62500000 --rebalance off --no-check > gpurun_out/bench3_n2_blobs_off.json 2>/dev/null
for w in rebal2 off; do grep "^{" gpurun_out/bench3_n2_blobs_$w.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$w', d['ms_per_step'], d.get('check'), d['config'].get('rebalance'), d['config']['ranks']['arrivals'])"; done
