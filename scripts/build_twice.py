import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if os.environ.get("WITH_TORCH"):
    import torch; torch.cuda.set_device(0); torch.zeros(1, device="cuda")
import parlayann_b200 as P
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
idx = P.GraphIndex(data_type="uint8", metric="Euclidian", max_degree=64, synthetic=dict(n=n, d=128, kind="uint8", seed=1234, stream=0, r=24, nc=256))
for i in range(3):
    t0 = time.time(); idx.build(R=64, L=128, alpha=1.2, num_passes=1, sort_neighbors=True); print(n, "build", i, round(time.time() - t0, 3), [round(x, 3) for x in idx.build_phase_seconds], flush=True)
