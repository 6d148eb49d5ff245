# C3 ground truth alone at N ranks (N = visible GPUs), shared bounds off and on
N=$(nvidia-smi -L | wc -l)
mkdir -p gpurun_out
for share in 0 1; do
  PANN_GT_SHARE=$share timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$share bench.py --workload groundtruth --gpus $N --steps 5 > gpurun_out/bench_gt_${N}gpu_share$share.json 2> gpurun_out/bench_gt_${N}gpu_share$share.err
  python - <<PY
import json
for l in open("gpurun_out/bench_gt_${N}gpu_share$share.json"):
    if l.startswith("{"):
        d=json.loads(l); print("N=$N share=$share wall %.2f ms local %.2f ms"%(d["value"]*1e3, d["detail"]["local_device_ms"]), d["detail"]["checksum"], [round(x*1e3,2) for x in d["detail"]["all_seconds"]])
PY
done
