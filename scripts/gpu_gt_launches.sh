mkdir -p gpurun_out
GT_RUNS=1 timeout 300 python scripts/gt_scale.py u8only > gpurun_out/gt_plain.log 2>&1 &&
GT_RUNS=1 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/gt_launches.csv timeout 600 python scripts/gt_scale.py u8only > gpurun_out/gt_ncu.log 2>&1
tail -2 gpurun_out/gt_ncu.log
