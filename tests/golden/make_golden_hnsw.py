"""Generates tests/golden/hnsw_*.npz: the reference's HNSW (oracle/_ref/ref_hnsw = algorithms/HNSW/HNSW.hpp with the one-line
unvisited_frontier fix of SURVEY.md §8c) built with PARLAY_NUM_THREADS=1 (deterministic levels: mt19937{0} in node order), the
model file's arrays, and HNSW::search results for a few (k, ef).  Run in the build container only."""
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
from parlayann_b200.hnsw_file import read_hnsw  # noqa: E402

EXT = {"uint8": "u8bin", "int8": "i8bin", "float": "fbin"}
# name, data_type, metric, n, d, nq, m, efc, m_l, alpha, generator r / nc
CASES = [
    ("hnsw_f32_mips", "float", "mips", 3000, 40, 60, 12, 40, 0.5, 1.0, 32, 1),
    ("hnsw_f32_l2", "float", "Euclidian", 2500, 32, 50, 8, 32, 0.4, 1.1, 24, 256),
    ("hnsw_u8_l2", "uint8", "Euclidian", 2500, 64, 50, 10, 48, 0.36, 1.0, 24, 256),
]
SEARCHES = [(10, 10), (10, 50), (1, 20), (5, 100)]


def main():
    assert O.ref_hnsw_path(), "build oracle/_ref first (make -C oracle ref)"
    tmp = tempfile.mkdtemp(prefix="pann_golden_")
    for name, dt, metric, n, d, nq, m, efc, m_l, alpha, r, nc in CASES:
        base = O.gen_points(dt, n, d, seed=77, r=r, nc=nc)
        qs = O.gen_points(dt, nq, d, seed=77, stream=1, r=r, nc=nc)
        bp, qp, mp = (os.path.join(tmp, f"{name}.{x}.{EXT[dt]}") for x in ("base", "query", "model"))
        fileio.write_bin(bp, base); fileio.write_bin(qp, qs)
        common = ["-data_type", dt, "-dist_func", metric, "-base", bp]
        O.ref_hnsw_run(["build"] + common + ["-m", m, "-efc", efc, "-m_l", m_l, "-alpha", alpha, "-out", mp], threads=1)
        M = read_hnsw(mp)
        g = dict(params=np.array([n, d, nq, m, efc, 77, r, nc], dtype=np.int64), fparams=np.array([m_l, alpha], dtype=np.float32),
                 level=M.level, counts=M.counts.astype(np.uint32), edges=M.edges, entrance=M.entrance, code=np.array([M.code_U, M.size_node], dtype=np.uint64))
        for si, (k, ef) in enumerate(SEARCHES):
            out = os.path.join(tmp, f"{name}.s{si}")
            O.ref_hnsw_run(["search"] + common + ["-model", mp, "-query", qp, "-k", k, "-ef", ef, "-out", out])
            res = O.parse_hnsw_results(out + ".res", nq)
            g[f"s{si}_cnt"] = np.array([len(a) for a, _ in res], dtype=np.uint32)
            g[f"s{si}_ids"] = np.concatenate([a for a, _ in res])
            g[f"s{si}_dists"] = np.concatenate([b for _, b in res])
            g[f"s{si}_cmps"] = np.fromfile(out + ".cmps", dtype=np.uint32)
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **g)
        print(name, "ok: levels", np.bincount(M.level), "entrances", M.entrance, "edges", len(M.edges))


if __name__ == "__main__":
    main()
