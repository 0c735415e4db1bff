#!/bin/bash
# round-2 call AA: hexkey - 3 x 8-bit histogram fast path (always on), blocks per SM 8 / 6 / 5
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py::test_oracle_at_size -m gpu -x -q > gpurun_out/pytest_hexkey.log 2>&1
echo "parity rc=$?"; tail -3 gpurun_out/pytest_hexkey.log
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --no-cpu --no-files"
for m in 8 6 5; do
  PCB_ENABLE_TUNING=1 PCB_HEXKEY_MINB=$m timeout 300 $B > gpurun_out/hexkey_m$m.json 2>/dev/null
  grep "^{" gpurun_out/hexkey_m$m.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']; print('minb $m', d['ms_per_step'], d['config']['step_ms'], {n: k[n]['ms_per_step'] for n in ('hexkey','radix_onesweep','contest')})"
done
for w in cube blobs; do timeout 300 $B --workload $w > gpurun_out/hexkey_$w.json 2>/dev/null; grep "^{" gpurun_out/hexkey_$w.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']; print('$w', d['ms_per_step'], d['config']['step_ms'], {n: k[n]['ms_per_step'] for n in ('hexkey',)})"; done
