mkdir -p gpurun_out
REPS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:pann_gt_shrink --launch-skip 3 -c 1 -f -o gpurun_out/r2_shrink_full python scripts/gt_shard_diag.py > gpurun_out/ncu_shrink.log 2>&1; tail -2 gpurun_out/ncu_shrink.log
