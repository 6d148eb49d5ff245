"""C3: brute-force ground truth k=100 over synthetic 10M x 96 float (Deep-shaped), base rows sharded over the ranks
(torchrun), NCCL all-gather of the local top-k + K5 merge on every rank.  Prints one JSON line on rank 0."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import parlayann_b200 as P
from parlayann_b200.sharded import shard_bounds, groundtruth_sharded

n = int(os.environ.get("GT_N", 10_000_000)); d = 96; nq = 10_000; k = 100
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
lo, hi = shard_bounds(n, rank, world)
# the generator is counter based: rows lo..hi of the global set are produced directly on this rank's GPU
idx = P.GraphIndex(data_type="float", metric="Euclidian", max_degree=4, device=lr,
                   synthetic=dict(n=hi - lo, d=d, kind="float_norm", seed=1234, stream=0, row0=lo))
idx.synthetic_queries(nq, kind="float_norm", seed=1234, stream=1)
times = []
for it in range(4):
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    gi, gd = groundtruth_sharded(idx, None, k=k, base_offset=lo)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    times.append(float(t.item()))
if rank == 0:
    best = min(times[1:])
    print(json.dumps({"what": f"ground truth {nq} x {n} x {d} f32, k={k}, base sharded x{world}", "n_gpus": world, "seconds": best,
                      "all": times, "tflops_algorithmic": 2.0 * nq * n * d / best / 1e12, "local_device_ms": idx.last_timing()[0],
                      "checksum": int(gi.astype(np.int64).sum() % 1000003), "dist_checksum": float(gd.astype(np.float64).sum())}), flush=True)
if world > 1:
    dist.barrier(); dist.destroy_process_group()
