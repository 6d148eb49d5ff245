# launch list (durations only) of one Vamana build, after a plain run of the same command
mkdir -p gpurun_out
N=${N:-10000000}
PANN_BUILD_VERBOSE=${VERBOSE:-0} timeout 600 python scripts/build_only.py $N > gpurun_out/build_plain.log 2>&1 && tail -3 gpurun_out/build_plain.log &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/build_launches_$N.csv timeout 900 python scripts/build_only.py $N > gpurun_out/build_ncu.log 2>&1
tail -2 gpurun_out/build_ncu.log
