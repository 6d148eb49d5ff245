"""float rows: build + search rates on a SIFT-shaped synthetic set (C1's shape at a larger n)"""
import os, sys, time, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parlayann_b200 as P
n = int(os.environ.get("N", "1000000")); d = int(os.environ.get("D", "128")); metric = os.environ.get("METRIC", "Euclidian")
idx = P.GraphIndex(data_type="float", metric=metric, max_degree=64, device=0, synthetic=dict(n=n, d=d, kind="float", seed=1234, stream=0, r=24, nc=256))
t0 = time.time(); idx.build(R=64, L=128, alpha=1.2 if metric == "Euclidian" else 1.0, num_passes=1, sort_neighbors=True); print("build seconds", round(time.time() - t0, 3), idx.build_phase_seconds)
idx.synthetic_queries(10000, kind="float", seed=1234, stream=1)
idx.groundtruth_resident(k=100); print("gt ms", idx.last_timing()[0])
gt_ids, gt_d = idx.fetch_groundtruth()
for beam in (10, 12, 16, 24, 32, 64):
    idx.search_resident(k=10, beam=beam, cut=1.35)
    ms = []
    for _ in range(5):
        idx.search_resident(k=10, beam=beam, cut=1.35); ms.append(idx.last_timing()[0])
    ids, dists, vis, cmps = idx.fetch_results()
    truth = (np.arange(100)[None, :] < 10) | (gt_d == gt_d[:, 9:10])
    rec = ((ids[:, :, None] == gt_ids[:, None, :]) & truth[:, None, :]).any(axis=2).mean()
    print(f"beam {beam}: recall {rec:.4f} cmps {cmps.mean():.0f}  {min(ms):.3f} ms / 10K q = {10000 / min(ms) / 1e3:.2f} M QPS  bytes/s {idx.last_bytes() / min(ms) / 1e6:.0f} GB/s")
idx.synthetic_queries(200000, kind="float", seed=1234, stream=7)
idx.search_resident(k=10, beam=12, cut=1.35); idx.search_resident(k=10, beam=12, cut=1.35)
print("200K queries beam 12:", round(idx.last_timing()[0], 2), "ms =", round(200000 / idx.last_timing()[0] / 1e3, 2), "M QPS,", round(idx.last_bytes() / idx.last_timing()[0] / 1e6), "GB/s")
