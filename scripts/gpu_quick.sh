#!/bin/bash
# parity tests + device bench (no e2e, no cpu); prints a compact summary
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --no-cpu --e2e-steps 0 --steps 10 "$@" 2>gpurun_out/bench.err > gpurun_out/bench.json; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['config']['step_ms'][:4])
print({k:v['ms_per_step'] for k,v in d['roofline']['kernels'].items()}, 'sum', round(sum(v['ms_per_step'] for v in d['roofline']['kernels'].values()),2))
PY
