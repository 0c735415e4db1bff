#!/bin/bash
# parity (full suite) with the halo contest forced, then default; bench terrain / cube / blobs default vs halo shapes
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
PCBTEST_ALLOW_TUNING=1 PCB_ENABLE_TUNING=1 PCB_CONTEST_HALO=1 PCB_OCT_3PASS=1 timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/pytest_halo.log 2>&1; echo "halo rc=$?" >> gpurun_out/pytest_halo.log
tail -5 gpurun_out/pytest_halo.log
B="python bench.py --no-cpu --no-files --e2e-steps 0"
for w in terrain cube blobs; do
  timeout 600 $B --workload $w > gpurun_out/benchf_${w}_base.json 2>/dev/null
  for v in 1 2 3; do
    PCB_ENABLE_TUNING=1 PCB_CONTEST_HALO=$v timeout 600 $B --workload $w > gpurun_out/benchf_${w}_halo$v.json 2>/dev/null
  done
  PCB_ENABLE_TUNING=1 PCB_OCT_3PASS=1 timeout 600 $B --workload $w > gpurun_out/benchf_${w}_oct3.json 2>/dev/null
  PCB_ENABLE_TUNING=1 PCB_OCT_3PASS=1 PCB_CONTEST_HALO=1 timeout 600 $B --workload $w > gpurun_out/benchf_${w}_oct3halo1.json 2>/dev/null
done
timeout 600 python bench.py --no-cpu --e2e-steps 2 > gpurun_out/benchf_full.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/benchf_full.json'));print('e2e_files',d.get('e2e_files'))"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/benchf_*.json')):
    try:
        d=json.load(open(f)); k=d['roofline']['kernels']
        print(f.split('benchf_')[1][:-5], d['ms_per_step'], {n:k[n]['ms_per_step'] for n in ('radix_onesweep','contest','hexkey','loser_emit','oct_scan','filter') if n in k})
    except Exception as e: print(f, 'ERR', e)
PY
