# one GPU call: smoke + GPU tests + bench (both arms) + ONE ncu pass (launch list of the bench's timed region)
set -x
mkdir -p gpurun_out
nvidia-smi -L; nproc; lscpu | grep "Model name"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
python -m pytest tests -x -q -m gpu 2>&1 | tail -5
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -c 1500 gpurun_out/bench_ref.json
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 3500 gpurun_out/bench.json
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --profile > gpurun_out/ncu_launch.log 2>&1
ls -la gpurun_out | head -30
