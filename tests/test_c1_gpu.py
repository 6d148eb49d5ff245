"""BASELINE.json configs[0] (C1) at its full size: Vamana build + beam search over a synthetic SIFT-shaped 100K x 128
float set, R = 64, L = 128 — the UNMODIFIED reference on the box's host cores (oracle/_ref/ref_driver: knn_index::build_index,
qsearchAll) against the CUDA path through the C ABI.  Float rows are summed in the reference's own element order, so the
graph and the search results are compared for identity, not for recall.  Timings of both sides are printed (-s) and written
to gpurun_out/c1_float_100k.json when that directory exists."""
import json
import os
import re
import time

import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c1_build_and_search_match_reference_at_full_size(tmp_path):
    if not O.ref_driver_path():
        pytest.skip("oracle/_ref/ref_driver was not built (needs /root/reference at build time)")
    import parlayann_b200 as P
    from parlayann_b200 import fileio
    n, d, nq, R, L = 100000, 128, 10000, 64, 128
    base = O.gen_points("float", n, d, seed=11)
    qs = O.gen_points("float", nq, d, seed=11, stream=1)
    bp, qp, out = str(tmp_path / "b.fbin"), str(tmp_path / "q.fbin"), str(tmp_path / "ref")
    fileio.write_bin(bp, base); fileio.write_bin(qp, qs)
    common = ["-data_type", "float", "-dist_func", "Euclidian", "-base", bp]
    txt = O.ref_run(["build"] + common + ["-R", R, "-L", L, "-alpha", 1.2, "-passes", 1, "-sort", 1, "-out", out])
    ref_build_s = float(re.search(r"REF_BUILD_SECONDS (\S+)", txt).group(1))
    g_ref = fileio.read_graph(out + ".graph")

    idx = P.GraphIndex(base, None, "float", "Euclidian", max_degree=R)
    idx.build(R=R, L=L, alpha=1.2, num_passes=1)          # first call pays the one-time allocations
    t0 = time.time(); idx.build(R=R, L=L, alpha=1.2, num_passes=1); gpu_build_s = time.time() - t0
    g = idx.graph()
    live = np.arange(g.shape[1])[None, :] <= g[:, :1]      # slot 0 = degree; the graph file drops the slots beyond it
    assert np.array_equal(g[:, 0], g_ref[:, 0]) and np.array_equal(g[live], g_ref[live])

    rec = {"config": "C1: 100000 x 128 f32, R=64 L=128 alpha=1.2, 10000 queries, k=10", "host_threads": O.nthreads(),
           "build_seconds": {"reference_cpu": ref_build_s, "gpu": gpu_build_s}, "search": []}
    for beam in (10, 20, 50, 128):
        txt = O.ref_run(["search"] + common + ["-graph", out + ".graph", "-query", qp, "-k", 10, "-beam", beam, "-reps", 3, "-out", out + f".s{beam}"])
        ref_qps = float(re.search(r"REF_SEARCH_QPS (\S+)", txt).group(1))
        want = np.fromfile(out + f".s{beam}.ids", dtype=np.uint32).reshape(nq, 10)
        idx.upload_queries(qs)
        idx.search_resident(k=10, beam=beam, cut=1.35)
        ms = []
        for _ in range(5):
            idx.search_resident(k=10, beam=beam, cut=1.35); ms.append(idx.last_timing()[0])
        ids, dists, vis, cmps = idx.fetch_results()
        assert np.array_equal(ids, want), f"beam {beam}"
        assert np.array_equal(vis, np.fromfile(out + f".s{beam}.visited", dtype=np.uint32))
        assert np.array_equal(cmps, np.fromfile(out + f".s{beam}.distances", dtype=np.uint32))
        t0 = time.time(); ids2, _, _, _ = idx.search(qs, k=10, beam=beam, cut=1.35); e2e_s = time.time() - t0
        assert np.array_equal(ids2, want)
        rec["search"].append({"beam": beam, "reference_cpu_qps": ref_qps, "gpu_device_qps": nq / (min(ms) * 1e-3), "gpu_host_buffers_qps": nq / e2e_s})
    print(json.dumps(rec))
    gd = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(gd):
        with open(os.path.join(gd, "c1_float_100k.json"), "w") as f:
            json.dump(rec, f, indent=1)
