#!/bin/bash
# round-2 call L: octant scan with the in-tile ranks overlapped with the look-back - parity, A/B, ncu of both
mkdir -p gpurun_out
export PCBTEST_ALLOW_TUNING=1
PCB_ENABLE_TUNING=1 PCB_OCT_VARIANT=3 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py::test_oracle_at_size -m gpu -x -q > gpurun_out/pytest_octov.log 2>&1
echo "ov parity rc=$?"; tail -3 gpurun_out/pytest_octov.log
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --no-cpu --no-files"
for w in terrain cube; do
  for v in 0 3 4 6; do
    PCB_ENABLE_TUNING=1 PCB_OCT_VARIANT=$v timeout 300 $B --workload $w > gpurun_out/octov_${w}_v$v.json 2>/dev/null
    grep "^{" gpurun_out/octov_${w}_v$v.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']; print('$w v$v', d['ms_per_step'], d['config']['step_ms'], 'oct', k['oct_scan'])"
  done
done
P="python bench.py --steps 1 --warmup 1 --e2e-steps 0 --no-cpu --no-files --no-clocks"
for v in 0 3; do
  PCB_ENABLE_TUNING=1 PCB_OCT_VARIANT=$v timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_octscan -s 5 -c 5 -o gpurun_out/r02_octscan_l$v $P > gpurun_out/ncu_full_oct$v.log 2>&1
  python scripts/ncu_summary.py gpurun_out/r02_octscan_l$v.ncu-rep > gpurun_out/r02_ncu_full_octscan_l$v.txt 2>&1
  rm -f gpurun_out/r02_octscan_l$v.ncu-rep
done
grep -E "Kernel Name|gpu__time_duration|l1tex__throughput|warps_active|barrier|issue_active|inst_executed" gpurun_out/r02_ncu_full_octscan_l0.txt | head -14
grep -E "Kernel Name|gpu__time_duration|l1tex__throughput|warps_active|barrier|issue_active|inst_executed" gpurun_out/r02_ncu_full_octscan_l3.txt | head -14
