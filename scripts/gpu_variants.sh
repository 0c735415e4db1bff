#!/bin/bash
# usage: gpu_variants.sh <variant>[+t][:ENV=val,...]  (+t: run the GPU tests with it first) ...   device-only bench per library variant (variants/libpcb_<v>.so);
# variants whose name ends in "i" run the instrumentation script instead
mkdir -p gpurun_out
cp point-cloud_b200/libpcb.so /tmp/libpcb_orig.so
for spec in "$@"; do
  v=${spec%%:*}; envs=""; [[ "$spec" == *:* ]] && envs=$(echo "${spec#*:}" | tr ',' ' ')
  tst=0; [[ "$v" == *+t ]] && { tst=1; v=${v%+t}; }
  cp point-cloud_b200/variants/libpcb_$v.so point-cloud_b200/libpcb.so
  if [ $tst = 1 ]; then env $envs timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3; fi
  if [[ "$v" == *i ]]; then
    echo "== $spec"; env $envs timeout 200 python scripts/gpu_radix_instr.py 2>&1 | tail -4
  else
    env $envs timeout 300 python bench.py --no-cpu --e2e-steps 0 --steps 6 --no-clocks 2>gpurun_out/bench.err > gpurun_out/bench_$v.json || tail -3 gpurun_out/bench.err
    python - "$spec" gpurun_out/bench_$v.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
k={k:v['ms_per_step'] for k,v in d['roofline']['kernels'].items()}
print('==', sys.argv[1], '|', d['ms_per_step'], {a:k[a] for a in k if a in ('contest','oct_scan','radix_onesweep','hexkey','loser_emit','l0_split') or a.startswith('radix_p')})
PY
  fi
done
cp /tmp/libpcb_orig.so point-cloud_b200/libpcb.so
