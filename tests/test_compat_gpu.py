"""A script written against the reference's Python surface (python/scripts/sift_test.py: `import _ParlayANNpy as pann; import
wrapper as wp`; build -> load -> batch_search_from_string -> check_recall) runs unchanged on parlayann_b200/compat."""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r'''
import _ParlayANNpy as pann
import wrapper as wp
import sys, time
D = sys.argv[1]
metric = "Euclidian"
wp.build_vamana_index(metric, "uint8", D + "/base.u8bin", D + "/graph_32", 32, 64, 1.2, False)
Index = wp.load_index(metric, "uint8", D + "/base.u8bin", D + "/graph_32")
for Q in [14, 23]:
    start = time.time()
    neighbors, distances = Index.batch_search_from_string(D + "/query.u8bin", 10, Q, True, 1000)
    end = time.time()
    print(neighbors.size)
    print("QPS: ", f'{neighbors.shape[0]/(end - start):.6f}')
    Index.check_recall(D + "/query.u8bin", D + "/gt.ibin", neighbors, 10)
ids = Index.single_search(wp.np.zeros(128, dtype="uint8") if hasattr(wp, "np") else __import__("numpy").zeros(128, dtype="uint8"), 10, 20, False, 1000)
assert ids.shape == (10,)
# HNSW through the same surface (module.cpp:123-126; wrapper forwards R, L, alpha, two_pass into graph_degree, efc, m_l, alpha)
wp.build_hnsw_index(metric, "uint8", D + "/base.u8bin", D + "/model.hnsw", 16, 64, 0.36, 1.0)
H = wp.load_index(metric, "uint8", D + "/base.u8bin", D + "/model.hnsw", True)
n2, d2 = H.batch_search_from_string(D + "/query.u8bin", 10, 40, False, -1)
H.check_recall(D + "/query.u8bin", D + "/gt.ibin", n2, 10)
print("DONE", pann.defaults.GRAPH_DEGREE, pann.UInt8EuclidianIndex.__name__)
'''


def test_reference_style_script(tmp_path, dataset):
    from parlayann_b200 import fileio
    ds = dataset("uint8", n=20000, d=128, nq=500)
    fileio.write_bin(str(tmp_path / "base.u8bin"), ds.base)
    fileio.write_bin(str(tmp_path / "query.u8bin"), ds.queries)
    gi, gd = O.groundtruth(ds.base, ds.queries, "uint8", k=100)
    fileio.write_ibin(str(tmp_path / "gt.ibin"), gi, gd)
    (tmp_path / "script.py").write_text(SCRIPT)
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([ROOT, os.path.join(ROOT, "parlayann_b200", "compat")]))
    r = subprocess.run([sys.executable, str(tmp_path / "script.py"), str(tmp_path)], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rec = [float(l.split(":")[1]) for l in r.stdout.splitlines() if l.startswith("Recall:")]
    assert len(rec) == 3 and min(rec) > 0.9, r.stdout
    assert "DONE 64 UInt8EuclidianIndex" in r.stdout
    # the graph file the script's build wrote is the reference's format and the oracle's graph
    g = fileio.read_graph(str(tmp_path / "graph_32"))
    ref, _, _ = O.vamana_build(ds.base, "uint8", R=32, L=64, alpha=float(np.float32(1.2)), sort_neighbors=True)
    m = np.arange(32)[None, :] < ref[:, 0][:, None]
    assert np.array_equal(g[:, 0], ref[:, 0]) and np.array_equal(np.sort(np.where(m, g[:, 1:], 0), axis=1), np.sort(np.where(m, ref[:, 1:], 0), axis=1))
