# ground-truth legs + shared bounds: tests, then C3 at N ranks with and without sharing
N=$(nvidia-smi -L | wc -l)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_groundtruth_gpu.py tests/test_sharded_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -5
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -2
for share in 1 0; do
  PANN_GT_SHARE=$share timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$share bench.py --workload groundtruth --gpus $N --steps 4 > gpurun_out/bench_gt_${N}gpu_share$share.json 2> gpurun_out/bench_gt_${N}gpu_share$share.err
  tail -c 900 gpurun_out/bench_gt_${N}gpu_share$share.json; tail -3 gpurun_out/bench_gt_${N}gpu_share$share.err
done
