"""Generates tests/golden/random_starts.npz: the reference's beamSearchRandom (algorithms/utils/beamSearch.h:309-351)
run by oracle/_ref/ref_driver on the u8_l2 golden case — per-query random start points, ids and counters.
Run in the build container only:  python tests/golden/make_golden_random.py"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as O  # noqa: E402
from parlayann_b200 import fileio  # noqa: E402


def main():
    assert O.ref_driver_path(), "build oracle/_ref first (make -C oracle ref)"
    tmp = tempfile.mkdtemp(prefix="pann_golden_")
    n, d, nq, R, L = 3000, 64, 200, 16, 32
    base = O.gen_points("uint8", n, d, seed=77)
    qs = O.gen_points("uint8", nq, d, seed=77, stream=1)
    bp, qp = os.path.join(tmp, "b.u8bin"), os.path.join(tmp, "q.u8bin")
    fileio.write_bin(bp, base); fileio.write_bin(qp, qs)
    common = ["-data_type", "uint8", "-dist_func", "Euclidian", "-base", bp]
    out = os.path.join(tmp, "r")
    O.ref_run(["build"] + common + ["-R", R, "-L", L, "-alpha", 1.2, "-sort", 0, "-out", out])
    g = {"params": np.array([n, d, nq, R, L, 77], dtype=np.int64), "graph": fileio.read_graph(out + ".graph")}
    for si, (k, beam) in enumerate([(10, 10), (10, 24), (5, 64)]):
        O.ref_run(["random"] + common + ["-graph", out + ".graph", "-query", qp, "-k", k, "-beam", beam, "-out", out + f".{si}"])
        g[f"r{si}_ids"] = np.fromfile(out + f".{si}.ids", dtype=np.uint32).reshape(nq, k)
        g[f"r{si}_visited"] = np.fromfile(out + f".{si}.visited", dtype=np.uint32)
        g[f"r{si}_distances"] = np.fromfile(out + f".{si}.distances", dtype=np.uint32)
        g[f"r{si}_kb"] = np.array([k, beam], dtype=np.int64)
        st = np.fromfile(out + f".{si}.starts", dtype=np.uint32)
        assert "starts" not in g or np.array_equal(g["starts"], st)
        g["starts"] = st
    # the start-point generator alone at a few (nq, n): dis(gen[i]) over graphs of other sizes
    np.savez_compressed(os.path.join(HERE, "random_starts.npz"), **g)
    print("random_starts ok", g["starts"][:8])


if __name__ == "__main__":
    main()
