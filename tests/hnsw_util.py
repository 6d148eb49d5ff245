"""Loader for tests/golden/hnsw_*.npz (outputs of the reference's HNSW; made by tests/golden/make_golden_hnsw.py)."""
import os

import numpy as np

import oracle_lib as O
from golden_util import GOLDEN

HNSW_CASES = {"hnsw_f32_mips": ("float", "mips"), "hnsw_f32_l2": ("float", "Euclidian"), "hnsw_u8_l2": ("uint8", "Euclidian")}
HNSW_SEARCHES = [(10, 10), (10, 50), (1, 20), (5, 100)]


class HnswGolden:
    def __init__(self, name):
        from parlayann_b200.hnsw_file import HnswModel
        self.name = name
        self.data_type, self.metric = HNSW_CASES[name]
        z = self.z = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.n, self.d, self.nq, self.m, self.efc, seed, r, nc = (int(x) for x in z["params"])
        self.m_l, self.alpha = (float(x) for x in z["fparams"])
        self.base = O.gen_points(self.data_type, self.n, self.d, seed=seed, r=r, nc=nc)
        self.queries = O.gen_points(self.data_type, self.nq, self.d, seed=seed, stream=1, r=r, nc=nc)
        self.model = HnswModel(self.d, self.m_l, self.m, self.efc, self.alpha, z["level"], z["counts"].astype(np.uint64), z["edges"], z["entrance"],
                               int(z["code"][0]), int(z["code"][1]))

    def search_result(self, si):
        cnt = self.z[f"s{si}_cnt"]
        off = np.concatenate([[0], np.cumsum(cnt.astype(np.int64))])
        ids, dists = self.z[f"s{si}_ids"], self.z[f"s{si}_dists"]
        return [(ids[off[i]:off[i + 1]], dists[off[i]:off[i + 1]]) for i in range(self.nq)], self.z[f"s{si}_cmps"]


def same_model(a, b):
    return (np.array_equal(a.level, b.level) and np.array_equal(a.entrance, b.entrance) and np.array_equal(a.counts, b.counts)
            and np.array_equal(a.edges, b.edges))


def check_search(got, golden_rows, golden_cmps, exact_ties):
    """got = (ids [nq, ocap], dists, cnt, cmps).  Entries tied in distance may come in any order in the reference (std::sort by
    distance only, HNSW.hpp:1596): with exact_ties=False rows are compared as (distance-sorted) sets per distance value."""
    ids, dists, cnt, cmps = got
    assert np.array_equal(cmps, golden_cmps), "distance comparison counts differ"
    for i, (ri, rd) in enumerate(golden_rows):
        assert cnt[i] == len(ri), (i, cnt[i], len(ri))
        c = min(len(ri), ids.shape[1])
        assert np.array_equal(dists[i, :c], rd[:c]), i
        if exact_ties:
            assert np.array_equal(ids[i, :c], ri[:c]), i
        elif c == len(ri):
            assert sorted(zip(dists[i, :c].tolist(), ids[i, :c].tolist())) == sorted(zip(rd.tolist(), ri.tolist())), i


def model_list_mismatches(a, b):
    """number of (node, level) neighbour lists that differ as SETS (the models must agree on levels and list sizes)"""
    assert np.array_equal(a.level, b.level) and np.array_equal(a.counts, b.counts)
    off = np.concatenate([[0], np.cumsum(a.counts.astype(np.int64))])
    bad = 0
    for i in np.nonzero(np.add.reduceat((a.edges != b.edges).astype(np.int64), off[:-1][a.counts > 0]) > 0)[0]:
        j = np.nonzero(a.counts > 0)[0][i]
        bad += set(a.edges[off[j]:off[j + 1]].tolist()) != set(b.edges[off[j]:off[j + 1]].tolist())
    return bad
