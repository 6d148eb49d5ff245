"""C5 (BASELINE configs[4]): HNSW build + search, synthetic 1M x 200 float inner product (Text2Image-shaped: r = 32, no clusters).
Prints build seconds (phases), search QPS / recall@10 for a few ef against exact ground truth (K4)."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import parlayann_b200 as P
from parlayann_b200.hnsw import HNSW
from parlayann_b200.harness import recall_with_ties

n = int(os.environ.get("N", 1000000)); d = 200; nq = 10000
m = int(os.environ.get("M", 32)); efc = int(os.environ.get("EFC", 128)); m_l = float(os.environ.get("ML", 0.36)); alpha = float(os.environ.get("ALPHA", 1.0))
h = HNSW(None, "float", "mips", m=m, efc=efc, m_l=m_l, alpha=alpha, synthetic=dict(n=n, d=d, kind="float", seed=1234, stream=0, r=32, nc=1))
t0 = time.time(); sec = h.build(); wall = time.time() - t0
info = h.info()
M = h.model()
print(json.dumps({"build_wall_s": wall, "phases(total,search,prune,reverse)": sec, "levels": np.bincount(M.level).tolist(), "entrances": len(M.entrance),
                  "avg_degree_l0": float(M.counts[np.concatenate([[0], np.cumsum(M.level.astype(np.int64) + 1)])[:-1]].mean()), "launches": h.index.last_timing()[1]}))
h.index.synthetic_queries(nq, kind="float", seed=1234, stream=1, r=32, nc=1)
qs = h.index.download_queries()
h.index.groundtruth_resident(k=100)
gi, gd = h.index.fetch_groundtruth()
for ef in (10, 20, 40, 80, 160):
    for _ in range(2):
        t0 = time.time(); ids, dists, cnt, cmps = h.search(qs, k=10, ef=ef); dt = time.time() - t0
    print(json.dumps({"ef": ef, "recall10": recall_with_ties(ids, gi.astype(np.uint32), gd, 10), "qps_e2e": nq / dt, "device_ms": h.index.last_timing()[0], "avg_cmps": float(cmps.mean())}))
