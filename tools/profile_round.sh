#!/bin/bash
# Captures of a round (run on the GPU box, outputs under gpurun_out/): the bench line, its ncu launch list, and
# `ncu --set full` reports (kept as raw CSV pages: the reports themselves exceed what travels back) of the hot kernels on C2, C3 and the D7 shard of 8 images.  Summaries are made on the
# build machine with tools/ncu_summary.py / tools/launch_list_summary.py and committed under profiles/.
set -u
T=${1:-r5}
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || tail -c 400 gpurun_out/${T}_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2>/dev/null
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-configs --no-shard-part > gpurun_out/${T}_ncu_bench.log 2>&1
M=smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum
timeout 900 ncu --set full --metrics $M --clock-control none -k regex:k7_ -s 7 -c 25 -o /tmp/${T}_c2 -f \
    python tools/prof_kernels.py c2_effdet_d0_512 --reps 2 --stages module_fwd_bwd,detect > /dev/null 2>&1
ncu -i /tmp/${T}_c2.ncu-rep --page raw --csv > gpurun_out/${T}_c2_raw.csv 2>/dev/null
timeout 900 ncu --set full --metrics $M --clock-control none -k regex:k7_ -s 7 -c 25 -o /tmp/${T}_c3 -f \
    python tools/prof_kernels.py c3_ksevendet_thermal_512x640 --reps 2 --stages module_fwd_bwd,detect > /dev/null 2>&1
ncu -i /tmp/${T}_c3.ncu-rep --page raw --csv > gpurun_out/${T}_c3_raw.csv 2>/dev/null
timeout 900 ncu --set full --metrics $M --clock-control none -k regex:k7_ -s 12 -c 35 -o /tmp/${T}_c5b8 -f \
    python tools/prof_kernels.py c5_effdet_d7_1536 --batch 8 --reps 2 --stages module_fwd_bwd,detect > /dev/null 2>&1
ncu -i /tmp/${T}_c5b8.ncu-rep --page raw --csv > gpurun_out/${T}_c5b8_raw.csv 2>/dev/null
for w in c1_retinanet_r50_512 c2_effdet_d0_512 c3_ksevendet_thermal_512x640 c4_effdet_d3_896 c5_effdet_d7_1536 "c5_effdet_d7_1536 --batch 8" "c4_effdet_d3_896 --batch 16"; do
  python tools/prof_kernels.py $w --reps 20 --stages match_zero,loss_fwd,loss_fwd_bwd,module_fwd,module_fwd_bwd,detect
done > gpurun_out/${T}_stage_times_all_configs.txt 2>&1
ls -la gpurun_out | tail -12
