"""one shard of C3 at N=8 (1.25M x 96 f32, 10 000 queries, k=100): device time of K4 as a function of the seed / growth"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import parlayann_b200 as P
n = int(os.environ.get("N", "1250000"))
idx = P.GraphIndex(data_type="float", metric="Euclidian", max_degree=1, synthetic=dict(n=n, d=96, kind="float_norm", seed=1234, stream=0))
idx.synthetic_queries(10000, kind="float_norm", seed=1234, stream=1)
for seed in os.environ.get("SEEDS", "512").split(","):
    os.environ["PANN_GT_SEED"] = seed
    ms = []
    for _ in range(int(os.environ.get("REPS", "4"))):
        idx.groundtruth_resident(k=100)
        ms.append(idx.last_timing())
    print(f"n={n} seed={seed}: " + " ".join(f"{m:.2f}ms/{l}" for m, l in ms), "mma launches", idx.groundtruth_stats(), flush=True)
