#!/bin/bash
# round-2 call Y: verification of the final tree (onesweep with the digit width compiled in) on one GPU
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_smoke_y.txt 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke_y.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_y.txt 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_y.txt
timeout 900 python bench.py --no-cpu > gpurun_out/r02_bench_y_terrain.json 2>gpurun_out/bench_y.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_y.err
for w in cube blobs city; do timeout 600 python bench.py --no-cpu --e2e-steps 0 --no-files --workload $w > gpurun_out/r02_bench_y_$w.json 2>/dev/null; done
for w in terrain cube blobs city; do grep "^{" gpurun_out/r02_bench_y_$w.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']; print('$w', d['value'], d['ms_per_step'], d['config']['step_ms'], (d.get('e2e') or {}).get('value'), k['kernel'], k['frac'], k['kernel_sum'], {n: v['ms_per_step'] for n, v in list(k['kernels'].items())[:5]})"; done
P="python bench.py --steps 1 --warmup 1 --e2e-steps 0 --no-cpu --no-files --no-clocks"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^k_' -c 600 --csv --log-file gpurun_out/r02_launches_y.csv $P > gpurun_out/ncu_list.log 2>&1
python scripts/ncu_launches.py gpurun_out/r02_launches_y.csv > gpurun_out/r02_ncu_launches_y.txt; head -9 gpurun_out/r02_ncu_launches_y.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_onesweep -s 30 -c 6 -o gpurun_out/r02_onesweep_y $P > gpurun_out/ncu_full_y.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02_onesweep_y.ncu-rep > gpurun_out/r02_ncu_full_onesweep_y.txt 2>&1
rm -f gpurun_out/r02_onesweep_y.ncu-rep
grep -E "Kernel Name|gpu__time_duration|issue_active|inst_executed|=> traffic" gpurun_out/r02_ncu_full_onesweep_y.txt | head -12 | cut -c1-140
