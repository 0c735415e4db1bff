#!/bin/bash
# round-2 call S: ncu --set full of one window of the second C2 conversion (all top kernels), final tree
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --e2e-steps 0 --no-cpu --no-files --no-clocks"
timeout 300 $CMD > gpurun_out/plain_s.log 2>&1 || { echo "plain failed"; exit 1; }
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_onesweep|k_chain_scan|k_octscan|k_loser_emit|k_hexkey|k_winner_emit|k_filter$|k_l0_split' -s 70 -c 22 -o gpurun_out/r02_final_window $CMD > gpurun_out/ncu_full_s.log 2>&1
echo "ncu rc=$?"
python scripts/ncu_summary.py gpurun_out/r02_final_window.ncu-rep gpurun_out/r02_traffic_final.json > gpurun_out/r02_ncu_full_final.txt 2>&1
grep -E "Kernel Name|gpu__time_duration|=> traffic" gpurun_out/r02_ncu_full_final.txt | cut -c1-150
cat gpurun_out/r02_traffic_final.json
rm -f gpurun_out/r02_final_window.ncu-rep
