mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-c3 --no-c5 --profile"
$B > gpurun_out/plain.log 2>&1 || { tail -20 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2_launches_bench_final.csv $B > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:pann_beam_search_small -c 1 -f -o gpurun_out/r2_k1s_final_full $B > gpurun_out/ncu_k1s.log 2>&1; tail -2 gpurun_out/ncu_k1s.log
