# multi-GPU round: NCCL tests + the bench under torchrun at N ranks (N = number of visible GPUs)
N=$(nvidia-smi -L | wc -l)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sharded_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 --no-c5 --no-cpu-baseline --no-c3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
tail -c 3000 gpurun_out/bench_${N}gpu.json; tail -5 gpurun_out/bench_${N}gpu.err
for share in 1 0; do
  PANN_GT_SHARE=$share timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$share bench.py --workload groundtruth --gpus $N --steps 4 > gpurun_out/bench_gt_${N}gpu_share$share.json 2> gpurun_out/bench_gt_${N}gpu_share$share.err
  tail -c 700 gpurun_out/bench_gt_${N}gpu_share$share.json; tail -2 gpurun_out/bench_gt_${N}gpu_share$share.err
done
