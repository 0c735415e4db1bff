#!/bin/bash
# round-2 call AI: ncu launch list + full capture of one unfiltered level on the FINAL tree
mkdir -p gpurun_out
P="python bench.py --steps 1 --warmup 1 --e2e-steps 0 --no-cpu --no-files --no-clocks"
timeout 300 $P > gpurun_out/plain_ai.log 2>&1 || { echo "plain failed"; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^k_' -c 600 --csv --log-file gpurun_out/r02_launches_final.csv $P > gpurun_out/ncu_list.log 2>&1
python scripts/ncu_launches.py gpurun_out/r02_launches_final.csv > gpurun_out/r02_ncu_launches_final.txt; head -12 gpurun_out/r02_ncu_launches_final.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_onesweep|k_chain_scan|k_octscan|k_loser_emit|k_hexkey|k_winner_emit' -s 82 -c 10 -o gpurun_out/r02_final_level $P > gpurun_out/ncu_full_ai.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02_final_level.ncu-rep gpurun_out/r02_traffic_final_level.json > gpurun_out/r02_ncu_full_final_level.txt 2>&1
grep -E "Kernel Name|gpu__time_duration|=> traffic" gpurun_out/r02_ncu_full_final_level.txt | cut -c1-130
rm -f gpurun_out/r02_final_level.ncu-rep
