#!/usr/bin/env python
"""One block of key metrics per DISTINCT kernel (its last launch) from `ncu -i X.ncu-rep --page raw --csv`.
  tools/ncu_last_launch_summary.py raw.csv "header line 1" ["header line 2"] > profiles/NAME.txt"""
import csv
import sys

KEEP = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'launch__shared_mem_per_block_dynamic', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__sass_thread_inst_executed_op_fadd_pred_on.sum', 'smsp__sass_thread_inst_executed_op_fmul_pred_on.sum',
        'smsp__sass_thread_inst_executed_op_ffma_pred_on.sum', 'sm__inst_executed_pipe_fp64.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio']


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    name = hdr.index('Kernel Name')
    last = {}
    for r in rows[2:]:
        last[r[name].split('(')[0]] = r
    for line in sys.argv[2:]:
        print(line)
    for r in last.values():
        print('---')
        for k in KEEP:
            if k in hdr:
                i = hdr.index(k)
                print(f"{k} [{units[i]}] = {r[i]}")


if __name__ == "__main__":
    main()
