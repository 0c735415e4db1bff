#!/bin/bash
# round-2 call V: three and five waves in the end-to-end path
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --e2e-steps 8 --no-cpu --no-files"
for wp in 25000000 34000000 20000000 25000000 34000000; do
  PCB_ENABLE_TUNING=1 PCB_WAVE_POINTS=$wp timeout 300 $B > gpurun_out/waves2_$wp.json 2>/dev/null
  grep "^{" gpurun_out/waves2_$wp.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); e=d['e2e']; print('wave_points $wp', 'e2e', e['value'], e['ms_per_step'], e.get('phases_ms'), 'device-only', d['ms_per_step'])"
done
