#!/bin/bash
# round-2 call AC: stream cache + pool-allocated staging buffer (open / close without driver locks) - full suite, bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_ac.txt 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_ac.txt
for i in 1 2; do
  timeout 600 python bench.py --no-cpu --no-files > gpurun_out/r02_bench_ac$i.json 2>/dev/null
  grep "^{" gpurun_out/r02_bench_ac$i.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); h=d['config']['step_host']; print('run $i', d['value'], d['ms_per_step'], d['config']['step_ms'], 'open/close max', max(x[0] for x in h['open_finish_close_ms']), max(x[2] for x in h['open_finish_close_ms']), 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e'].get('phases_ms'))"
done
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
