"""A/B of the float-row staging in the general kernel K1 (PANN_NO_STAGE=1 turns it off): build and wide-beam search"""
import os, sys, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parlayann_b200 as P
n = int(os.environ.get("N", "1000000")); d = int(os.environ.get("D", "128"))
for rep in range(2):
    idx = P.GraphIndex(data_type="float", metric="Euclidian", max_degree=64, device=0, synthetic=dict(n=n, d=d, kind="float", seed=1234, stream=0, r=24, nc=256))
    t0 = time.time(); idx.build(R=64, L=128, alpha=1.2, num_passes=1, sort_neighbors=True); print("NO_STAGE", os.environ.get("PANN_NO_STAGE", "0"), "build seconds", round(time.time() - t0, 3), [round(x, 3) for x in idx.build_phase_seconds])
idx.synthetic_queries(10000, kind="float", seed=1234, stream=1)
for beam in (12, 48, 64, 100):
    idx.search_resident(k=10, beam=beam, cut=1.35)
    ms = []
    for _ in range(5):
        idx.search_resident(k=10, beam=beam, cut=1.35); ms.append(idx.last_timing()[0])
    print(f"  beam {beam}: {min(ms):.3f} ms / 10K q")
