#!/bin/bash
# round-2 call AB: hexkey - key assembly with a per-cell base, floor-converts, 32-bit indices
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py::test_oracle_at_size -m gpu -x -q > gpurun_out/pytest_hexkey2.log 2>&1
echo "parity rc=$?"; tail -3 gpurun_out/pytest_hexkey2.log
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --no-cpu --no-files"
for w in terrain blobs cube; do timeout 300 $B --workload $w > gpurun_out/hexkey2_$w.json 2>/dev/null; grep "^{" gpurun_out/hexkey2_$w.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']; print('$w', d['ms_per_step'], d['config']['step_ms'], {n: k[n]['ms_per_step'] for n in ('hexkey','filter','loser_emit')})"; done
