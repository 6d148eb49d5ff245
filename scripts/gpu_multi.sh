# multi-GPU round: NCCL tests + the bench under torchrun at N ranks (N = number of visible GPUs)
N=$(nvidia-smi -L | wc -l)
mkdir -p gpurun_out
python -m pytest tests/test_sharded_gpu.py -x -q -m gpu 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
tail -c 3000 gpurun_out/bench_${N}gpu.json; tail -5 gpurun_out/bench_${N}gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --workload groundtruth --gpus $N --steps 3 > gpurun_out/bench_gt_${N}gpu.json 2> gpurun_out/bench_gt_${N}gpu.err
tail -c 1200 gpurun_out/bench_gt_${N}gpu.json
