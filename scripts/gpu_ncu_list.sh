#!/bin/bash
# usage: gpu_ncu_list.sh <tag> [bench args...]   launch list (durations) of the library's kernels
TAG=$1; shift 1
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --e2e-steps 0 --no-cpu $@"
timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^k_' -c 2000 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
