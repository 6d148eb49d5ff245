# round-2 profiles: launch list of the bench's timed region + ncu --set full of K1s, K3 (both modes, 10M build), K4 (tcgen05)
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-c3 --no-c5 --profile"
$B > gpurun_out/plain.log 2>&1 || { tail -20 gpurun_out/plain.log; exit 1; }
tail -c 600 gpurun_out/plain.log
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2_launches_bench.csv $B > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:pann_beam_search_small -c 1 -f -o gpurun_out/r2_k1s_full $B > gpurun_out/ncu_k1s.log 2>&1; tail -2 gpurun_out/ncu_k1s.log
# K3 in a 10M build: skip to a full-size batch (4 prune launches per batch: visited lists, its long-list launch, reverse groups, its long-list launch)
python scripts/build_only.py 10000000 > gpurun_out/build_plain.log 2>&1; tail -1 gpurun_out/build_plain.log
ncu --set full --clock-control none --import-source on -k regex:pann_prune_kernel --launch-skip 240 -c 4 -f -o gpurun_out/r2_k3_full python scripts/build_only.py 10000000 > gpurun_out/ncu_k3.log 2>&1; tail -2 gpurun_out/ncu_k3.log
# K4: the largest chunk of a 10 000 x 10M x 128 u8 ground truth
ncu --set full --clock-control none --import-source on -k regex:pann_gt_mma --launch-skip 8 -c 1 -f -o gpurun_out/r2_k4_full GT_RUNS=1 python scripts/gt_scale.py big > gpurun_out/ncu_k4.log 2>&1; tail -2 gpurun_out/ncu_k4.log
ls -la gpurun_out/*.ncu-rep
