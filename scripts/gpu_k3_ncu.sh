mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pann_prune_kernel --launch-skip 242 -c 1 -f -o gpurun_out/r2_k3_rev_full python scripts/build_only.py 10000000 > gpurun_out/ncu_k3.log 2>&1; tail -2 gpurun_out/ncu_k3.log
