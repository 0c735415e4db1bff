#!/bin/bash
# round-2 call Z: are slow steps host preemption?  default bench line three times with the per-step host diagnostics
mkdir -p gpurun_out
for i in 1 2 3; do
  timeout 300 python bench.py --no-cpu --e2e-steps 0 --no-files --steps 10 > gpurun_out/r02_bench_z$i.json 2>/dev/null
  grep "^{" gpurun_out/r02_bench_z$i.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('run $i', d['value'], d['ms_per_step'], d['config']['step_ms'], d['config']['step_host'])"
done
nproc; cat /proc/loadavg; ps -eo pcpu,comm --sort=-pcpu | head -8
