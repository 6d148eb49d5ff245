"""ground-truth timing at scale: K4 v2 (tcgen05) vs v1; synthetic data generated on the device"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import parlayann_b200 as P
cfgs = [("uint8", "Euclidian", "uint8", 1000000, 128, 10000, 100), ("float", "Euclidian", "float_norm", 1000000, 96, 10000, 100)]
if len(sys.argv) > 1 and sys.argv[1] == "big":
    cfgs = [("uint8", "Euclidian", "uint8", 10000000, 128, 10000, 100), ("float", "Euclidian", "float_norm", 10000000, 96, 10000, 100)]
if len(sys.argv) > 1 and sys.argv[1] == "u8only":
    cfgs = cfgs[:1]
for dt, metric, kind, n, d, nq, k in cfgs:
    idx = P.GraphIndex(data_type=dt, metric=metric, max_degree=4, synthetic=dict(n=n, d=d, kind=kind, seed=1234, stream=0))
    idx.synthetic_queries(nq, kind=kind, seed=1234, stream=1)
    res = {}
    for v1 in ("0", "0", "1")[:int(os.environ.get("GT_RUNS", "3"))]:
        os.environ["PANN_GT_V1"] = v1
        idx.groundtruth_resident(k=k)
        ms, nl = idx.last_timing()
        res[v1] = idx.fetch_groundtruth()
        flops = 2.0 * nq * n * d
        print(f"{dt} n={n} d={d} nq={nq} k={k} {'v1' if v1 == '1' else 'v2'}: {ms:.2f} ms, {nl} launches, {flops / ms / 1e9:.1f} T{'OP' if dt != 'float' else 'FLOP'}/s algorithmic", flush=True)
    if "1" in res:
        print("   v2 == v1:", np.array_equal(res["0"][1], res["1"][1]), np.array_equal(res["0"][0], res["1"][0]))
    idx.close()
