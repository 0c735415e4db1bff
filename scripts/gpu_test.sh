#!/bin/bash
# parity tests + short bench (+ trace)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
if grep -q "pytest rc=0" gpurun_out/pytest_gpu.log; then
  PCB_TRACE=1 timeout 300 python bench.py --steps 1 --warmup 1 --e2e-steps 0 --no-cpu > gpurun_out/trace.log 2>&1
  grep "\[pcb\]" gpurun_out/trace.log | tail -45
  timeout 600 python bench.py "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
  cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
fi
