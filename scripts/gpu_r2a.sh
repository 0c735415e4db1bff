#!/bin/bash
# round 2, call A: parity tests (incl. the at-size oracle test) + the new bench workloads
mkdir -p gpurun_out
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt; nvidia-smi -L >> gpurun_out/host.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -22 gpurun_out/pytest_gpu.log
for w in terrain cube merge blobs city; do
  extra="--no-cpu --no-files"
  [ $w = terrain ] && extra="--cpu-scale 0.05"
  timeout 600 python bench.py --workload $w $extra > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
  cat gpurun_out/bench_$w.json | cut -c1-1800; tail -3 gpurun_out/bench_$w.err
done
timeout 300 python bench.py --workload cube --points 1000000 --no-cpu --no-files > gpurun_out/bench_cube1m.json 2> gpurun_out/bench_cube1m.err
cat gpurun_out/bench_cube1m.json | cut -c1-600
