#!/bin/bash
# usage: gpu_round_profile.sh <tag>   full bench line, ncu launch list of the same command, ncu --set full of one level
TAG=$1
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench_$TAG.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2>> gpurun_out/bench_$TAG.err; echo "ref rc=$?"
bash scripts/gpu_ncu_list.sh $TAG
bash scripts/gpu_ncu_full.sh $TAG 'k_onesweep|k_chain_scan|k_octscan|k_loser_emit|k_hexkey|k_winner_emit' 27 9
