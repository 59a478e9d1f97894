timeout 300 ncu --set full --clock-control none --import-source on -k regex:k7_nms_brute_kernel -s 1 -c 1 -o gpurun_out/r5g_brute_c2 -f python tools/prof_kernels.py c2_effdet_d0_512 --reps 2 --stages detect > /dev/null 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k7_nms_pairs_kernel -s 1 -c 1 -o gpurun_out/r5g_pairs_c3 -f python tools/prof_kernels.py c3_ksevendet_thermal_512x640 --reps 2 --stages detect > /dev/null 2>&1
ls -la gpurun_out/
