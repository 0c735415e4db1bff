#!/bin/bash
# round-2 call N: verification of the final tree on one GPU - smoke, full GPU suite, default bench line (with the CPU legs),
# the reference arm, the other workloads
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_smoke_n.txt 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke_n.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_n.txt 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_n.txt
timeout 1200 python bench.py > gpurun_out/r02_bench_n_terrain.json 2>gpurun_out/bench_n.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_n.err
timeout 1200 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r02_bench_n_ref.json 2>gpurun_out/bench_n_ref.err; echo "ref rc=$?"; tail -c 600 gpurun_out/r02_bench_n_ref.json
for w in cube blobs city merge; do timeout 600 python bench.py --no-cpu --e2e-steps 0 --no-files --workload $w > gpurun_out/r02_bench_n_$w.json 2>/dev/null; done
for w in terrain cube blobs city merge; do grep "^{" gpurun_out/r02_bench_n_$w.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']; print('$w', d['value'], d['ms_per_step'], d['config']['step_ms'], (d.get('e2e') or {}).get('value'), (d.get('e2e_files') or ''), k['kernel'], k['frac'], k['kernel_sum'], d.get('cpu_baseline'))"; done
