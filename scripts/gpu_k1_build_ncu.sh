mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pann_beam_search_kernel --launch-skip 40 -c 1 -f -o gpurun_out/r2_k1_build_full python scripts/build_only.py 10000000 > gpurun_out/ncu_k1b.log 2>&1; tail -2 gpurun_out/ncu_k1b.log
