#!/bin/bash
# round-2 call H: octant-scan variant 3/4 (register diet) - parity, then A/B on three workloads; merge loader check;
# ncu capture of the contest on the skewed cloud
mkdir -p gpurun_out
export PCBTEST_ALLOW_TUNING=1
for v in 3 4; do
  PCB_ENABLE_TUNING=1 PCB_OCT_VARIANT=$v timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py::test_oracle_at_size -m gpu -x -q > gpurun_out/pytest_oct$v.log 2>&1
  echo "oct variant $v parity rc=$?"; tail -3 gpurun_out/pytest_oct$v.log
done
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --no-cpu --no-files"
for w in terrain cube blobs; do
  for v in 0 3 4; do
    PCB_ENABLE_TUNING=1 PCB_OCT_VARIANT=$v timeout 300 $B --workload $w > gpurun_out/oct_${w}_v$v.json 2>/dev/null
    grep "^{" gpurun_out/oct_${w}_v$v.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']; print('$w v$v', d['ms_per_step'], d['config']['step_ms'], 'oct', k['oct_scan'])"
  done
done
# merge loader (node files parsed as views)
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_cli.py -m gpu -x -q -k "merge or reingest or own" > gpurun_out/pytest_merge.log 2>&1; echo "merge tests rc=$?"; tail -3 gpurun_out/pytest_merge.log
timeout 600 python bench.py --steps 3 --warmup 1 --no-cpu --workload merge > gpurun_out/bench_merge.json 2>gpurun_out/bench_merge.err; tail -c 1500 gpurun_out/bench_merge.json
# contest + octant scan on the skewed cloud, second conversion
timeout 1200 ncu --set full --clock-control none --kernel-name-base demangled -k regex:'ContestOp|k_octscan' -s 24 -c 24 -o gpurun_out/r02_blobs_scans \
  python bench.py --steps 1 --warmup 1 --e2e-steps 0 --no-cpu --no-files --no-clocks --workload blobs > gpurun_out/ncu_full_blobs.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02_blobs_scans.ncu-rep > gpurun_out/r02_ncu_full_blobs_contest_octscan.txt 2>&1
rm -f gpurun_out/r02_blobs_scans.ncu-rep
wc -l gpurun_out/r02_ncu_full_blobs_contest_octscan.txt
