#!/bin/bash
# round-2 profiles: launch list of the C2 step + full captures (onesweep on C2; contest on C4 and C1 shapes)
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --e2e-steps 0 --no-cpu --no-files --no-clocks"
timeout 300 $B > gpurun_out/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^k_' -c 600 --csv --log-file gpurun_out/r02_launches.csv $B > gpurun_out/ncu_list.log 2>&1
python scripts/ncu_launches.py gpurun_out/r02_launches.csv > gpurun_out/r02_ncu_launches.txt; head -25 gpurun_out/r02_ncu_launches.txt
# dominant kernel, second conversion (skip the first conversion's 23 launches), levels 2-3
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_onesweep -s 30 -c 6 -o gpurun_out/r02_onesweep $B > gpurun_out/ncu_full1.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02_onesweep.ncu-rep gpurun_out/r02_traffic_onesweep.json > gpurun_out/r02_ncu_full_onesweep.txt 2>&1
# contest + octant scan on the skewed cloud (all levels of one conversion) and on the cube's chain levels
for w in blobs cube; do
  timeout 1200 ncu --set full --clock-control none -k regex:'ContestOp|k_octscan' -s 24 -c 28 -o gpurun_out/r02_${w}_scans $B --workload $w > gpurun_out/ncu_full_$w.log 2>&1
  python scripts/ncu_summary.py gpurun_out/r02_${w}_scans.ncu-rep > gpurun_out/r02_ncu_full_${w}_contest_octscan.txt 2>&1
  rm -f gpurun_out/r02_${w}_scans.ncu-rep
done
ls -la gpurun_out | tail -12
# compute-sanitizer logs of small conversions through every sort path (memcheck always; racecheck bounded by a timeout)
timeout 600 compute-sanitizer --tool memcheck --log-file gpurun_out/r02_sanitizer_memcheck.txt python scripts/gpu_sanitize.py > gpurun_out/sanitize_mem.log 2>&1; echo "memcheck rc=$?"; tail -3 gpurun_out/r02_sanitizer_memcheck.txt
timeout 900 compute-sanitizer --tool racecheck --log-file gpurun_out/r02_sanitizer_racecheck.txt python scripts/gpu_sanitize.py > gpurun_out/sanitize_race.log 2>&1; echo "racecheck rc=$?"; tail -3 gpurun_out/r02_sanitizer_racecheck.txt
