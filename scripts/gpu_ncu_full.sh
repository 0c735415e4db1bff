#!/bin/bash
# usage: gpu_ncu_full.sh <tag> <kernel regex> <skip> <count> [bench args...]   (one ncu run per GPU call)
TAG=$1; RE=$2; SKIP=$3; CNT=$4; shift 4
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --e2e-steps 0 --no-cpu $@"
timeout 300 $CMD > gpurun_out/plain2_$TAG.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$RE" -s $SKIP -c $CNT -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out | tail -4
