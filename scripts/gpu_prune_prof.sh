mkdir -p gpurun_out
timeout 600 python scripts/build_only.py 10000000 > gpurun_out/build_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pann_prune_kernel -s 120 -c 3 -f -o gpurun_out/prune_full timeout 900 python scripts/build_only.py 10000000 > gpurun_out/prune_ncu.log 2>&1
tail -3 gpurun_out/prune_ncu.log
