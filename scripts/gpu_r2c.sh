#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/pytest_noalloc_alone.log
tail -4 gpurun_out/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --cpu-scale 0.05 > gpurun_out/bench_terrain.json 2> gpurun_out/bench_terrain.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/bench_terrain.json
