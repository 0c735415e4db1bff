#!/bin/bash
# round-2 call X: onesweep with the digit width compiled in - parity, then A/B against the generic kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py::test_oracle_at_size -m gpu -x -q > gpurun_out/pytest_radix8.log 2>&1
echo "parity rc=$?"; tail -3 gpurun_out/pytest_radix8.log
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --no-cpu --no-files"
for w in terrain blobs; do
  for g in 0 1; do
    if [ $g = 1 ]; then PCB_ENABLE_TUNING=1 PCB_RADIX_GENERIC=1 timeout 300 $B --workload $w > gpurun_out/radix8_${w}_g$g.json 2>/dev/null; else timeout 300 $B --workload $w > gpurun_out/radix8_${w}_g$g.json 2>/dev/null; fi
    grep "^{" gpurun_out/radix8_${w}_g$g.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']; print('$w generic=$g', d['ms_per_step'], d['config']['step_ms'], {n: k[n]['ms_per_step'] for n in ('radix_onesweep','l0_split')}, d['roofline']['frac'])"
  done
done
