mkdir -p gpurun_out
GT_RUNS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:pann_gt_mma --launch-skip 8 -c 1 -f -o gpurun_out/r2_k4_full python scripts/gt_scale.py big > gpurun_out/ncu_k4.log 2>&1; tail -2 gpurun_out/ncu_k4.log
