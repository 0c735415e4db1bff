#!/bin/bash
# usage: gpu_sweep.sh "ENV=val ENV2=val" "ENV=val" ...   one device-only bench per environment set
mkdir -p gpurun_out
for cfg in "$@"; do
  env $cfg python bench.py --no-cpu --e2e-steps 0 --steps 6 --no-clocks 2>gpurun_out/bench.err > gpurun_out/bench.json || tail -3 gpurun_out/bench.err
  python - "$cfg" <<'PY'
import json,sys
d=json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
k={k:v['ms_per_step'] for k,v in d['roofline']['kernels'].items()}
print(sys.argv[1], '|', d['ms_per_step'], {a:k[a] for a in k if a in ('contest','oct_scan','radix_onesweep','hexkey','loser_emit') or a.startswith('radix_p')})
PY
done
