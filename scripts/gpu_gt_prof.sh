mkdir -p gpurun_out
timeout 300 python scripts/gt_scale.py u8only > gpurun_out/gt_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/gt_launches.csv timeout 600 python scripts/gt_scale.py u8only > gpurun_out/gt_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pann_gt_mma -s 2 -c 1 -f -o gpurun_out/gt_mma_full timeout 600 python scripts/gt_scale.py u8only > gpurun_out/gt_ncu_full.log 2>&1
tail -3 gpurun_out/gt_ncu_full.log
