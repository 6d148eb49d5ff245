import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import parlayann_b200 as P
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
idx = P.GraphIndex(data_type="uint8", metric="Euclidian", max_degree=64, synthetic=dict(n=n, d=128, kind="uint8", seed=1234, stream=0))
t0 = time.time(); idx.build(R=64, L=128, alpha=1.2, num_passes=1, sort_neighbors=os.environ.get('SORT', '0') == '1'); print("build seconds", time.time() - t0, idx.build_phase_seconds)
