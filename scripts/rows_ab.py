"""search rates for row shapes other than 128-byte integer rows (generic QR = 0 kernels); PANN_LIB_PATH selects the build"""
import os, sys, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parlayann_b200 as P
n = int(os.environ.get("N", "1000000"))
for dt, metric, d in (("int8", "mips", 200), ("uint8", "Euclidian", 96), ("uint8", "Euclidian", 256)):
    idx = P.GraphIndex(data_type=dt, metric=metric, max_degree=64, device=0, synthetic=dict(n=n, d=d, kind=dt, seed=1234, stream=0))
    t0 = time.time(); idx.build(R=64, L=128, alpha=1.2 if metric == "Euclidian" else 1.0, num_passes=1, sort_neighbors=True)
    print(dt, metric, d, "build seconds", round(time.time() - t0, 3), [round(x, 3) for x in idx.build_phase_seconds])
    idx.synthetic_queries(200000, kind=dt, seed=1234, stream=1)
    for beam in (12, 24, 64):
        idx.search_resident(k=10, beam=beam, cut=1.35)
        ms = []
        for _ in range(3):
            idx.search_resident(k=10, beam=beam, cut=1.35); ms.append(idx.last_timing()[0])
        print(f"  beam {beam}: {min(ms):.3f} ms / 200K q = {200000 / min(ms) / 1e3:.2f} M QPS, {idx.last_bytes() / min(ms) / 1e6:.0f} GB/s")
    del idx
