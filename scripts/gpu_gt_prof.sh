mkdir -p gpurun_out
GT_RUNS=1 timeout 300 python scripts/gt_scale.py u8only > gpurun_out/gt_plain.log 2>&1 &&
GT_RUNS=1 ncu --set full --clock-control none --import-source on -k regex:pann_gt_mma -s 5 -c 1 -f -o gpurun_out/gt_mma_full timeout 600 python scripts/gt_scale.py u8only > gpurun_out/gt_ncu_full.log 2>&1
tail -2 gpurun_out/gt_ncu_full.log
GT_RUNS=2 timeout 600 python scripts/gt_scale.py big 2>&1 | tee gpurun_out/gt_big.log
