"""Generates tests/golden/batch_insert.npz: the reference's knn_index::batch_insert (vamana/index.h:188-316) called directly three
times on one graph by oracle/_ref/ref_driver (`insert`): first half of the points in the given order, second half in random
order, a scattered subset again (alpha 1.0, max_fraction .05); the graph after every call and the accumulated statistics."""
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
    tmp = tempfile.mkdtemp(prefix="pann_golden_")
    n, d, R, L = 2600, 48, 12, 32
    base = O.gen_points("uint8", n, d, seed=77)
    bp = os.path.join(tmp, "b.u8bin")
    fileio.write_bin(bp, base)
    out = os.path.join(tmp, "ins")
    O.ref_run(["insert", "-data_type", "uint8", "-dist_func", "Euclidian", "-base", bp, "-R", R, "-L", L, "-alpha", 1.2, "-out", out])
    g = dict(params=np.array([n, d, R, L, 77], dtype=np.int64))
    for i in range(3):
        g[f"g{i}"] = fileio.read_graph(out + f".g{i}")
    g["bvisited"] = np.fromfile(out + ".bvisited", dtype=np.uint32)
    g["bdistances"] = np.fromfile(out + ".bdistances", dtype=np.uint32)
    np.savez_compressed(os.path.join(HERE, "batch_insert.npz"), **g)
    print("batch_insert ok", [int(g[f"g{i}"][:, 0].sum()) for i in range(3)])


if __name__ == "__main__":
    main()
