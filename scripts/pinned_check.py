import os, sys, time, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parlayann_b200 as P
from parlayann_b200.index import pinned_empty
from parlayann_b200.sharded import search_sharded
n, nq, k = 1000000, 10000, 10
idx = P.GraphIndex(data_type="uint8", metric="Euclidian", max_degree=64, device=0, synthetic=dict(n=n, d=128, kind="uint8", seed=1234, stream=0, r=24, nc=256))
idx.build(R=64, L=128, alpha=1.2)
idx.synthetic_queries(nq, kind="uint8", seed=1234, stream=1)
q = torch.from_numpy(idx.download_queries()).pin_memory().numpy()
def outs(kind):
    sh = [((nq, k), np.uint32), ((nq, k), np.float32), ((nq,), np.uint32), ((nq,), np.uint32)]
    if kind == "pageable": return tuple(np.empty(s, d) for s, d in sh)
    if kind == "torch": return tuple(torch.empty(s, dtype={np.uint32: torch.int32, np.float32: torch.float32}[d]).pin_memory().numpy().view(d) for s, d in sh)
    return tuple(pinned_empty(s, d) for s, d in sh)
for kind in ("pageable", "torch", "pann", "sharded"):
    o = None if kind == "sharded" else outs(kind)
    ts = []
    for r in range(12):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        res = search_sharded(idx, q, k=k, beam=16) if kind == "sharded" else idx.search(q, k=k, beam=16, out=o)
        ts.append((time.perf_counter() - t0) * 1e3)
    print(kind, " ".join(f"{t:.3f}" for t in ts), flush=True)
