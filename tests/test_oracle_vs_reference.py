"""Pins the checker: oracle/pann_oracle.cpp (our C++ restatement) against outputs of the UNMODIFIED
reference (tests/golden/*.npz, made by tests/golden/make_golden.py through oracle/_ref/ref_driver).
CPU only.  Integer element types: bit-exact.  Floats: the restatement reproduces g++'s sequential float
sum (oracle/Makefile: -ffp-contract=off), so it is checked bit-exact too — including rows whose length is not a multiple of
8 elements (f32_*_d100: the reference binary built here with its own flags, -O3 -march=native, produces the sequential
non-contracted sums there as well)."""
import numpy as np
import pytest

import oracle_lib as O
from golden_util import CASES, SEARCHES, golden, valid_graph, rows_equal_as_sets


@pytest.mark.parametrize("name", list(CASES))
def test_build_graph_matches_reference(name):
    g = golden(name)
    got, sv, sd = O.vamana_build(g.base, g.data_type, g.metric, R=g.R, L=g.L, alpha=1.2, passes=g.passes, sort_neighbors=False)
    assert np.array_equal(valid_graph(got), valid_graph(g.z["graph_unsorted"])), "unsorted graph differs from the reference"
    assert np.array_equal(sv, g.z["build_visited"]) and np.array_equal(sd, g.z["build_distances"])
    got_s, _, _ = O.vamana_build(g.base, g.data_type, g.metric, R=g.R, L=g.L, alpha=1.2, passes=g.passes, sort_neighbors=True)
    assert rows_equal_as_sets(got_s, g.z["graph_sorted"])
    assert np.array_equal(valid_graph(got_s), valid_graph(g.z["graph_sorted"])), "sorted graph differs (same libstdc++ expected)"


@pytest.mark.parametrize("name", list(CASES))
def test_search_matches_reference(name):
    g = golden(name)
    ran = 0
    for si, s in enumerate(SEARCHES):
        if f"s{si}_ids" not in g.z:
            continue
        r = O.search(g.base, g.graph, g.queries, g.data_type, g.metric, **s)
        assert np.array_equal(r["ids"], g.z[f"s{si}_ids"]), (name, s)
        assert np.array_equal(r["visited"], g.z[f"s{si}_visited"]), (name, s)
        assert np.array_equal(r["cmps"], g.z[f"s{si}_distances"]), (name, s)
        ran += 1
    assert ran >= 4


@pytest.mark.parametrize("name", list(CASES))
def test_full_lists_match_reference(name):
    g = golden(name)
    r = O.search(g.base, g.graph, g.queries, g.data_type, g.metric, k=0, beam=g.L, cut=0.0, full=True, vcap=2048)
    cnt = g.z["full_counts"]
    assert np.array_equal(r["nfr"], cnt[:, 0]) and np.array_equal(r["nvis"], cnt[:, 1]) and np.array_equal(r["cmps"], cnt[:, 2])
    fr = g.split_lists(cnt[:, 0], g.z["full_frontier"])
    vi = g.split_lists(cnt[:, 1], g.z["full_visited"])
    for i in range(g.nq):
        assert np.array_equal(r["fr_ids"][i, :cnt[i, 0]], fr[i][:, 0])
        assert np.array_equal(r["fr_dists"][i, :cnt[i, 0]].view(np.uint32), fr[i][:, 1])
        assert np.array_equal(r["vis_ids"][i, :cnt[i, 1]], vi[i][:, 0])
        assert np.array_equal(r["vis_dists"][i, :cnt[i, 1]].view(np.uint32), vi[i][:, 1])
    # build mode: the query is base row i (same_as skips it)
    selfs = np.arange(g.nq)
    r = O.search(g.base, g.graph, g.base[:g.nq], g.data_type, g.metric, k=0, beam=g.L, cut=0.0, selfs=selfs, full=True, vcap=2048)
    cnt = g.z["self_counts"]
    assert np.array_equal(r["nvis"], cnt[:, 1]) and np.array_equal(r["cmps"], cnt[:, 2])
    vi = g.split_lists(cnt[:, 1], g.z["self_visited"])
    for i in range(g.nq):
        assert np.array_equal(r["vis_ids"][i, :cnt[i, 1]], vi[i][:, 0])


@pytest.mark.parametrize("name", list(CASES))
def test_groundtruth_matches_reference(name):
    g = golden(name)
    ids, dists = O.groundtruth(g.base, g.queries, g.data_type, g.metric, k=g.gt_k)
    assert np.array_equal(dists, g.z["gt_dists"]), "sorted distance rows must be identical"
    # ids: identical as sets per tie group (the reference's sort is by distance only, not stable)
    ref = g.z["gt_ids"]
    for i in range(g.nq):
        if np.array_equal(ids[i], ref[i]):
            continue
        for dv in np.unique(dists[i]):
            m = dists[i] == dv
            if dv == dists[i, -1]:
                continue  # the boundary tie group may be cut differently
            assert set(ids[i][m].tolist()) == set(ref[i][m].tolist())


def test_random_permutation_matches_reference():
    z = np.load(golden("u8_l2").z.fid.name.replace("u8_l2.npz", "perms.npz"))
    for key in z.files:
        n = int(key.split("_")[1])
        assert np.array_equal(O.random_permutation(n), z[key]), n


def test_distance_known_answers():
    L = O.lib()
    import ctypes as C
    a = np.zeros(128, np.uint8); b = np.full(128, 255, np.uint8)
    assert L.oracle_distance(1, 0, a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), 128) == 8323200.0
    c = np.full(128, -128, np.int8)
    assert L.oracle_distance(2, 1, c.ctypes.data_as(C.c_void_p), c.ctypes.data_as(C.c_void_p), 128) == -2097152.0


def test_batch_schedule_known_answer():
    s = O.batch_schedule(100000)
    assert s[:4] == [(0, 1), (1, 3), (3, 7), (7, 15)] and s[10] == (1023, 2047) and s[11] == (2047, 4047)
    assert len(s) == 60 and s[-1][1] == 100000


def insert_steps(n):
    """the three batch_insert calls of ref_driver `insert` / tests/golden/make_golden_insert.py"""
    return [(np.arange(n // 2, dtype=np.uint32), 1.2, False, .02), (np.arange(n // 2, n, dtype=np.uint32), 1.2, True, .02),
            (((np.arange(n // 7, dtype=np.uint64) * 7 + 3) % n).astype(np.uint32), 1.0, False, .05)]


def test_batch_insert_matches_reference():
    """knn_index::batch_insert used directly (vamana/index.h:188-316): subsets, random order, another max_fraction, re-insertion"""
    import os
    from golden_util import GOLDEN, valid_graph
    z = np.load(os.path.join(GOLDEN, "batch_insert.npz"))
    n, d, R, L, seed = (int(x) for x in z["params"])
    base = O.gen_points("uint8", n, d, seed=seed)
    g = np.zeros((n, R + 1), np.uint32)
    sv = np.zeros(n, np.uint32); sd = np.zeros(n, np.uint32)
    for i, (ins, alpha, rnd, frac) in enumerate(insert_steps(n)):
        a, b = O.vamana_batch_insert(base, g, ins, "uint8", L=L, alpha=alpha, random_order=rnd, max_fraction=frac)
        sv += a; sd += b
        assert np.array_equal(valid_graph(g), valid_graph(z[f"g{i}"])), i
    assert np.array_equal(sv, z["bvisited"]) and np.array_equal(sd, z["bdistances"])
