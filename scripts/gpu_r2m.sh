#!/bin/bash
# round-2 call M: spin back-off in the look-backs (variant libraries built with -DPCB_SPIN_BACKOFF / -DPCB_RADIX_BACKOFF)
mkdir -p gpurun_out
cp point-cloud_b200/libpcb.so gpurun_out/libpcb_base.so
B="python bench.py --steps 5 --warmup 3 --e2e-steps 0 --no-cpu --no-files"
for lib in base s40 s100 s250 r40 r100s100; do
  if [ $lib = base ]; then cp gpurun_out/libpcb_base.so point-cloud_b200/libpcb.so; else cp point-cloud_b200/variants/libpcb_$lib.so point-cloud_b200/libpcb.so; fi
  for v in 0 3; do
    PCB_ENABLE_TUNING=1 PCB_OCT_VARIANT=$v timeout 300 $B > gpurun_out/backoff_${lib}_v$v.json 2>/dev/null
    grep "^{" gpurun_out/backoff_${lib}_v$v.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']; print('$lib v$v', d['ms_per_step'], d['config']['step_ms'], {n: k[n]['ms_per_step'] for n in ('radix_onesweep','contest','oct_scan','l0_split')})"
  done
done
cp gpurun_out/libpcb_base.so point-cloud_b200/libpcb.so; rm -f gpurun_out/libpcb_base.so
