#!/usr/bin/env python
"""Summarise an `ncu --page raw --csv` dump (one row per captured launch) into a markdown table for profiles/."""
import csv
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__warps_eligible.avg.per_cycle_active',
        'lts__t_bytes.sum', 'l1tex__t_bytes.sum']
STALL = 'smsp__average_warp'   # smsp__average_warps_issue_stalled_*_per_issue_active.ratio (ncu 2025: warp_latency...)


def main(path, stalls=True):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    kn = hdr.index('Kernel Name')
    for r in rows[2:]:
        print(f"\n## {r[kn][:110]}\n")
        print("| metric | value | unit |\n|---|---|---|")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f"| {w} | {r[i]} | {units[i]} |")
        if stalls:
            st = []
            for i, h in enumerate(hdr):
                if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('_per_issue_active.ratio') and 'not_issued' not in h:
                    try:
                        st.append((float(r[i].replace(',', '')), h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
                    except ValueError:
                        pass
            st.sort(reverse=True)
            if st:
                print("\ntop warp stall reasons (warps stalled per issue-active cycle): " + ", ".join(f"{n} {v:.2f}" for v, n in st[:6]))


if __name__ == "__main__":
    main(sys.argv[1])
