"""does anything known before the search predict how long a query's walk is? (a longest-first order would shorten a batch's tail)"""
import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parlayann_b200 as P
n = 10000000
idx = P.GraphIndex(data_type="uint8", metric="Euclidian", max_degree=64, device=0, synthetic=dict(n=n, d=128, kind="uint8", seed=1234, stream=0, r=24, nc=256))
idx.build(R=64, L=128, alpha=1.2)
idx.synthetic_queries(10000, kind="uint8", seed=1234, stream=1)
q = idx.download_queries().astype(np.int32)
idx.search_resident(k=10, beam=24, cut=1.35)
ids, dists, vis, cmps = idx.fetch_results()
start = idx.points()[0].astype(np.int32) if hasattr(idx, "points") else None
d0 = ((q - start[None, :]) ** 2).sum(1)
print("hops mean %.1f max %d" % (vis.mean(), vis.max()))
for name, x in (("dist to start", d0), ("dist to best result", dists[:, 0]), ("dist to 10th result", dists[:, 9]), ("query norm", (q ** 2).sum(1))):
    print(name, "corr with hops %.3f" % np.corrcoef(x, vis)[0, 1], " with cmps %.3f" % np.corrcoef(x, cmps)[0, 1])
