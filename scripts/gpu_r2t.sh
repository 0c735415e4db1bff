#!/bin/bash
# round-2 call T: number of waves in the end-to-end path (PCB_WAVE_POINTS), C2
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --e2e-steps 6 --no-cpu --no-files"
for wp in 0 50000000 25000000 12500000 8400000 6250000; do
  if [ $wp = 0 ]; then timeout 300 $B > gpurun_out/waves_$wp.json 2>/dev/null; else PCB_ENABLE_TUNING=1 PCB_WAVE_POINTS=$wp timeout 300 $B > gpurun_out/waves_$wp.json 2>/dev/null; fi
  grep "^{" gpurun_out/waves_$wp.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); e=d['e2e']; print('wave_points $wp', 'e2e', e['value'], e['ms_per_step'], e.get('phases_ms'), e.get('step_ms'))"
done
