# ncu full capture of one K1 launch of the bench workload (after a plain run of the same command)
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile ${BENCH_ARGS} > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:${KREGEX:-pann_beam_search} -c 1 -f -o gpurun_out/${OUT:-k1_full} \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile ${BENCH_ARGS} > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
