# full GPU suite (the no-alloc test runs last), then the K4 ncu capture
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
GT_RUNS=1 python scripts/gt_scale.py big > gpurun_out/gt_plain.log 2>&1; tail -2 gpurun_out/gt_plain.log
GT_RUNS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:pann_gt_mma --launch-skip 8 -c 1 -f -o gpurun_out/r2_k4_full python scripts/gt_scale.py big > gpurun_out/ncu_k4.log 2>&1; tail -2 gpurun_out/ncu_k4.log
ls -la gpurun_out/*.ncu-rep
