#!/usr/bin/env python
"""Summarise an .ncu-rep (one block per profiled launch) — run here, no GPU needed.
usage: scripts/ncu_summary.py gpurun_out/prof.ncu-rep [> profiles/...txt]"""
import csv
import subprocess
import sys

WANT = ['Kernel Name', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__shared_mem_per_block_dynamic', 'launch__shared_mem_per_block_static',
        'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_warps',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_membar_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum']


TRAFFIC = {}


def kernel_class(name):
    """ncu kernel name -> the class name the library's kernel statistics use (pcb_kernel_stat_get)."""
    table = [('k_onesweep', 'radix_onesweep'), ('ContestOp', 'contest'), ('k_octscan', 'oct_scan'),
             ('k_loser_emit', 'loser_emit'), ('k_winner_emit', 'winner_emit'), ('k_hexkey', 'hexkey'),
             ('CellScanOp', 'cell_scan'), ('L0HeadOp', 'l0_heads'), ('k_l0_gather', 'l0_gather'),
             ('k_l0_key', 'l0_key'), ('k_bbox_check', 'bbox_check')]
    for k, v in table:
        if k in name:
            return v
    return name.split('(')[0]


def main():
    rep = sys.argv[1]
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(w) for w in WANT if w in hdr]
    for r in rows[2:]:
        print('---')
        for i in idx:
            print(f'  {hdr[i]:78s} {r[i][:90]} {units[i]}')
        try:
            rd = float(r[hdr.index('dram__bytes_read.sum')].replace(',', ''))
            wr = float(r[hdr.index('dram__bytes_write.sum')].replace(',', ''))
            t = float(r[hdr.index('gpu__time_duration.sum')].replace(',', ''))
            mult = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
            ur, uw = units[hdr.index('dram__bytes_read.sum')], units[hdr.index('dram__bytes_write.sum')]
            ut = units[hdr.index('gpu__time_duration.sum')]
            tm = {'ns': 1e-9, 'us': 1e-6, 'ms': 1e-3, 's': 1}[ut]
            tot = rd * mult[ur] + wr * mult[uw]
            print(f'  => traffic {tot / 1e6:.1f} MB, {tot / (t * tm) / 1e9:.0f} GB/s DRAM')
            cls = kernel_class(r[hdr.index('Kernel Name')])
            TRAFFIC.setdefault(cls, []).append(tot)
        except Exception as e:  # noqa
            print('  => (no traffic summary)', e)


if __name__ == '__main__':
    main()
    if len(sys.argv) > 2:   # second argument: write {class: mean DRAM bytes per launch} for bench.py's roofline.traffic
        import json
        json.dump({k: round(sum(v) / len(v)) for k, v in TRAFFIC.items()}, open(sys.argv[2], 'w'), indent=1)
