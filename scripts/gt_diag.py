import os, sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import parlayann_b200 as P
import oracle_lib as O
for dt, kind, d in (("uint8","uint8",128), ("float","float_norm",96)):
    base = O.gen_points(kind, 100000, d); qs = O.gen_points(kind, 300, d, stream=1)
    idx = P.GraphIndex(base, None, dt, "Euclidian", max_degree=4)
    os.environ["PANN_GT_V1"] = "0"
    a = idx.groundtruth(qs, k=10); st = idx.groundtruth_stats(); ms = idx.last_timing()
    os.environ["PANN_GT_V1"] = "1"
    b = idx.groundtruth(qs, k=10)
    print(dt, "mma launches/fallbacks", st, "ms", ms, "dists equal", np.array_equal(a[1], b[1]), "ids equal", np.array_equal(a[0], b[0]), "rows differing", int((a[1] != b[1]).any(axis=1).sum()))
