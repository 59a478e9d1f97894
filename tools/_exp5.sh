timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -x --timeout 100 -k "nms or detect or postprocess or static" 2>&1 | tail -5
for w in "c2_effdet_d0_512" "c3_ksevendet_thermal_512x640" "c1_retinanet_r50_512" "c5_effdet_d7_1536 --batch 8" "c4_effdet_d3_896"; do
  timeout 120 python tools/prof_kernels.py $w --reps 20 --stages detect 2>&1 | tail -1
done
i=0
for w in "c2_effdet_d0_512" "c3_ksevendet_thermal_512x640" "c5_effdet_d7_1536 --batch 8"; do
 i=$((i+1))
 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k7_ -c 60 --csv --log-file gpurun_out/r5f_launches_$i.csv python tools/prof_kernels.py $w --reps 2 --stages detect > /dev/null 2>&1
done
