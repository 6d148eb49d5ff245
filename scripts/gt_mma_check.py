"""K4 v2 (tcgen05/TMA) against K4 v1 (CUDA cores, PANN_GT_V1=1) and the oracle; run on the GPU box."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O
import parlayann_b200 as P

cases = [("uint8", "Euclidian", "uint8", 20000, 128, 300, 100), ("uint8", "mips", "uint8", 20000, 128, 300, 10),
         ("int8", "Euclidian", "int8", 30011, 100, 257, 50), ("int8", "mips", "int8", 20000, 128, 100, 100),
         ("float", "Euclidian", "float_norm", 40000, 96, 300, 100), ("float", "mips", "float", 20000, 200, 130, 20),
         ("float", "Euclidian", "float", 20000, 128, 300, 100), ("uint8", "Euclidian", "uint8", 200000, 128, 1000, 100)]
if len(sys.argv) > 1:
    cases = cases[:int(sys.argv[1])]
for dt, metric, kind, n, d, nq, k in cases:
    base = O.gen_points(kind, n, d); qs = O.gen_points(kind, nq, d, stream=1)
    idx = P.GraphIndex(base, None, dt, metric, max_degree=4)
    os.environ["PANN_GT_V1"] = "0"
    t0 = time.time(); gi, gd = idx.groundtruth(qs, k=k); t1 = time.time(); ms2 = idx.last_timing()
    os.environ["PANN_GT_V1"] = "1"
    vi, vd = idx.groundtruth(qs, k=k); ms1 = idx.last_timing()
    ok_d = np.array_equal(gd, vd); ok_i = np.array_equal(gi, vi)
    msg = f"{dt:6s} {metric:9s} n={n} d={d} nq={nq} k={k}: v2==v1 dists {ok_d} ids {ok_i}  v2 {ms2[0]:.2f} ms ({ms2[1]} launches)  v1 {ms1[0]:.2f} ms"
    if n <= 40000:
        oi, od = O.groundtruth(base, qs, dt, metric, k=k)
        msg += f"  oracle dists {np.array_equal(gd, od)}"
    print(msg, flush=True)
    if not (ok_d and ok_i):
        bad = np.argwhere(gd != vd)
        print("  first mismatches", bad[:5], gd[bad[0][0]][:8] if len(bad) else "", vd[bad[0][0]][:8] if len(bad) else "")
        sys.exit(1)
print("gt_mma_check: all ok")
