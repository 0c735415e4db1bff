#!/bin/bash
# round-2 call J: octant scan v2 (packed codes) at 4-7 blocks per SM
mkdir -p gpurun_out
export PCBTEST_ALLOW_TUNING=1
for v in 6; do
  PCB_ENABLE_TUNING=1 PCB_OCT_VARIANT=$v timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_oct$v.log 2>&1
  echo "oct variant $v parity rc=$?"; tail -3 gpurun_out/pytest_oct$v.log
done
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --no-cpu --no-files"
for w in terrain cube; do
  for v in 0 5 6 7 8; do
    PCB_ENABLE_TUNING=1 PCB_OCT_VARIANT=$v timeout 300 $B --workload $w > gpurun_out/oct_${w}_v$v.json 2>/dev/null
    grep "^{" gpurun_out/oct_${w}_v$v.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']; print('$w v$v', d['ms_per_step'], d['config']['step_ms'], 'oct', k['oct_scan'])"
  done
done
