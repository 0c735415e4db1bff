#!/bin/bash
# round-2 call K: new default octant scan - full GPU suite, bench lines, merge loader A/B, ncu launch list + octscan capture
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_k.txt 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_k.txt
B="python bench.py --no-cpu"
timeout 600 $B > gpurun_out/r02_bench_k_terrain.json 2>gpurun_out/bench_k.err; echo "bench rc=$?"
for w in cube blobs city; do timeout 300 $B --e2e-steps 0 --no-files --workload $w > gpurun_out/r02_bench_k_$w.json 2>/dev/null; done
for w in terrain cube blobs city; do grep "^{" gpurun_out/r02_bench_k_$w.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']; print('$w', d['value'], d['ms_per_step'], (d.get('e2e') or {}).get('value'), k['kernel'], k['frac'], k['kernel_sum'], {n: v['ms_per_step'] for n, v in list(k['kernels'].items())[:6]})"; done
# merge loader A/B (same box, same files): new library, then the library of commit 8bec9e3
timeout 600 $B --steps 5 --warmup 2 --e2e-steps 0 --no-files --workload merge > gpurun_out/r02_bench_k_merge_new.json 2>/dev/null
cp point-cloud_b200/libpcb.so gpurun_out/libpcb_new.so; cp point-cloud_b200/variants/libpcb_oldloader.so point-cloud_b200/libpcb.so
timeout 600 $B --steps 5 --warmup 2 --e2e-steps 0 --no-files --workload merge > gpurun_out/r02_bench_k_merge_old.json 2>/dev/null
cp gpurun_out/libpcb_new.so point-cloud_b200/libpcb.so; rm -f gpurun_out/libpcb_new.so
for w in new old; do grep "^{" gpurun_out/r02_bench_k_merge_$w.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('merge $w', d['ms_per_step'], d['config']['step_ms'])"; done
# profiles
P="python bench.py --steps 1 --warmup 1 --e2e-steps 0 --no-cpu --no-files --no-clocks"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^k_' -c 600 --csv --log-file gpurun_out/r02_launches_k.csv $P > gpurun_out/ncu_list.log 2>&1
python scripts/ncu_launches.py gpurun_out/r02_launches_k.csv > gpurun_out/r02_ncu_launches_k.txt; head -12 gpurun_out/r02_ncu_launches_k.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_octscan -s 5 -c 5 -o gpurun_out/r02_octscan_k $P > gpurun_out/ncu_full_oct.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02_octscan_k.ncu-rep > gpurun_out/r02_ncu_full_octscan_k.txt 2>&1
ncu -i gpurun_out/r02_octscan_k.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
hdr=rows[0]
want=['l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum','l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum','l1tex__t_requests_pipe_lsu_mem_global_op_st.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','smsp__inst_executed.sum','sm__warps_active.avg.pct_of_peak_sustained_active','l1tex__throughput.avg.pct_of_peak_sustained_elapsed']
for r in rows[2:]:
    print({w: r[hdr.index(w)] for w in want if w in hdr})
" > gpurun_out/r02_octscan_k_l1.txt 2>&1
head -5 gpurun_out/r02_octscan_k_l1.txt
rm -f gpurun_out/r02_octscan_k.ncu-rep
