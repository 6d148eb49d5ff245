"""time of one search launch as a function of the batch size (fixed overhead vs per-query cost) for several PANN_K1S_TUNE values"""
import os, sys, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parlayann_b200 as P

n = int(os.environ.get("N", "10000000")); beam = int(os.environ.get("BEAM", "24"))
tunes = os.environ.get("TUNES", "auto,0,1,2").split(",")
idx = P.GraphIndex(data_type="uint8", metric="Euclidian", max_degree=64, device=0, synthetic=dict(n=n, d=128, kind="uint8", seed=1234, stream=0, r=24, nc=256))
idx.build(R=64, L=128, alpha=1.2, num_passes=1, sort_neighbors=True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ref = None
for nq in (1250, 2500, 5000, 10000, 100000, 1000000):
    idx.synthetic_queries(nq, kind="uint8", seed=1234, stream=1)
    line = f"nq {nq:7d}"
    for t in tunes:
        os.environ.pop("PANN_K1S_TUNE", None) if t == "auto" else os.environ.__setitem__("PANN_K1S_TUNE", t)
        for _ in range(2):
            idx.search_resident(k=10, beam=beam, cut=1.35)
        ms = []
        for _ in range(5):
            flush.zero_(); torch.cuda.synchronize()
            idx.search_resident(k=10, beam=beam, cut=1.35)
            ms.append(idx.last_timing()[0])
        line += f"  tune {t}: {sum(ms) / len(ms):.4f} ms"
        if nq == 10000:
            ids, _, vis, cmps = idx.fetch_results()
            if ref is None:
                ref = ids.copy()
                print("hops: avg %.1f p50 %d p99 %d max %d; cmps avg %.0f" % (vis.mean(), np.percentile(vis, 50), np.percentile(vis, 99), vis.max(), cmps.mean()))
            else:
                assert np.array_equal(ref, ids)
    print(line, flush=True)
