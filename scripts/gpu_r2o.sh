#!/bin/bash
# round-2 call O (--gpus 2): final tree on two GPUs - the 2-GPU tests, the C2 bench line with check
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 python -m pytest tests/test_gpu_p2p.py tests/test_gpu_shard.py -m gpu -x -q > gpurun_out/r02_pytest_p2p_n2_final.txt 2>&1; echo "p2p rc=$?"; tail -3 gpurun_out/r02_pytest_p2p_n2_final.txt
timeout 900 $TR bench.py --gpus 2 --no-cpu > gpurun_out/r02_bench_n2_final_terrain.json 2> gpurun_out/bench_n2.err; echo "terrain rc=$?"; tail -2 gpurun_out/bench_n2.err
grep "^{" gpurun_out/r02_bench_n2_final_terrain.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['config']['step_ms'], d.get('check'), (d.get('e2e') or {}).get('value'))"
