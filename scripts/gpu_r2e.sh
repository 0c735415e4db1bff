#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shard.py tests/test_cull.py -m gpu -x -q > gpurun_out/pytest_shard.log 2>&1; echo "shard rc=$?" >> gpurun_out/pytest_shard.log
tail -30 gpurun_out/pytest_shard.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
