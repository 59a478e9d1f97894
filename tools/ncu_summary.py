#!/usr/bin/env python
"""Key metrics per launch from `ncu -i X.ncu-rep --page raw --csv` output.
  tools/ncu_summary.py raw.csv > profiles/NAME.txt"""
import csv
import sys

KEEP = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'launch__shared_mem_per_block_dynamic', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'sm__inst_executed_pipe_fp64.sum',
        'smsp__sass_thread_inst_executed_op_fadd_pred_on.sum', 'smsp__sass_thread_inst_executed_op_fmul_pred_on.sum',
        'smsp__sass_thread_inst_executed_op_ffma_pred_on.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio']

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print('---')
    for k in KEEP:
        if k in hdr:
            i = hdr.index(k)
            print(f"{k} [{units[i]}] = {r[i]}")
    try:   # FP32 fraction: (FADD + FMUL + 2 FFMA) / duration against 148 SMs x 128 lanes x 2 x 1.965 GHz = 74.4 TFLOP/s
        g = lambda k: float(r[hdr.index(k)].replace(',', ''))
        fl = g('smsp__sass_thread_inst_executed_op_fadd_pred_on.sum') + g('smsp__sass_thread_inst_executed_op_fmul_pred_on.sum') + \
            2 * g('smsp__sass_thread_inst_executed_op_ffma_pred_on.sum')
        dur = g('gpu__time_duration.sum')
        dur_s = dur * (1e-6 if units[hdr.index('gpu__time_duration.sum')] in ('us', 'usecond') else 1e-9 if units[hdr.index('gpu__time_duration.sum')] in ('ns', 'nsecond') else 1e-3)
        print(f"fp32_flop = {fl:.0f}; fp32_frac_of_74.4_TFLOPs = {fl / dur_s / 74.4e12:.4f}")
    except Exception:
        pass
