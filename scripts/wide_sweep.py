"""K1s WIDE variant (both rounds' gathers in flight) against the default kernel as a function of the batch size"""
import os, sys, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parlayann_b200 as P
n = int(os.environ.get("N", "10000000")); beam = int(os.environ.get("BEAM", "24"))
idx = P.GraphIndex(data_type="uint8", metric="Euclidian", max_degree=64, device=0, synthetic=dict(n=n, d=128, kind="uint8", seed=1234, stream=0, r=24, nc=256))
idx.build(R=64, L=128, alpha=1.2, num_passes=1, sort_neighbors=True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for nq in [int(x) for x in os.environ.get("NQS", "1,32,500,1250,1776,2500,5000,10000,100000").split(",")]:
    idx.synthetic_queries(nq, kind="uint8", seed=1234, stream=1)
    line, ref = f"nq {nq:7d}", None
    for w in ("0", "1"):
        os.environ["PANN_K1S_WIDE"] = w
        for _ in range(2):
            idx.search_resident(k=10, beam=beam, cut=1.35)
        ms = []
        for _ in range(7):
            flush.zero_(); torch.cuda.synchronize()
            idx.search_resident(k=10, beam=beam, cut=1.35)
            ms.append(idx.last_timing()[0])
        res = idx.fetch_results()
        if ref is None: ref = res
        else: assert all(np.array_equal(a, b) for a, b in zip(ref, res)), "WIDE changed the results"
        line += f"   wide={w}: {np.median(ms):.4f} ms"
    print(line, flush=True)
