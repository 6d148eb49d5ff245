"""K1 (build mode) + K3 (robustPrune) + reverse-edge grouping = pann_vamana_build, through the C ABI.
Gates: with sort_neighbors=false the graph is byte-identical to the reference's (golden
fixtures) and to the oracle's at larger sizes; build statistics identical; with the final sort the rows are
identical too (rows without ties are rank-sorted on the device, rows with ties by a kernel that runs std::sort's algorithm)."""
import numpy as np
import pytest

import oracle_lib as O
from golden_util import CASES, INT_CASES, golden, valid_graph, rows_equal_as_sets

pytestmark = pytest.mark.gpu


def _build(base, data_type, metric, R, L, alpha=1.2, passes=1, sort_neighbors=False, stats=True):
    import parlayann_b200 as P
    idx = P.GraphIndex(base, None, data_type, metric, max_degree=R)
    st = idx.build(R=R, L=L, alpha=alpha, num_passes=passes, sort_neighbors=sort_neighbors, stats=stats)
    return idx, idx.graph(), st


@pytest.mark.parametrize("name", list(CASES))
def test_build_matches_reference_golden(name):
    g = golden(name)
    idx, got, (sv, sd) = _build(g.base, g.data_type, g.metric, g.R, g.L, passes=g.passes)
    assert np.array_equal(valid_graph(got), valid_graph(g.z["graph_unsorted"])), "graph differs from the reference's"
    assert np.array_equal(sv, g.z["build_visited"]), "build visited counters differ"
    assert np.array_equal(sd, g.z["build_distances"]), "build distance-comparison counters differ"
    _, got_s, _ = _build(g.base, g.data_type, g.metric, g.R, g.L, passes=g.passes, sort_neighbors=True, stats=False)
    assert np.array_equal(valid_graph(got_s), valid_graph(g.z["graph_sorted"])), "sorted graph differs from the reference's"


@pytest.mark.parametrize("data_type,metric,n,d,R,L,passes", [
    ("uint8", "Euclidian", 20000, 128, 32, 64, 1),
    ("uint8", "Euclidian", 6000, 128, 64, 128, 1),
    ("int8", "Euclidian", 6000, 100, 24, 48, 2),
    ("uint8", "mips", 5000, 64, 16, 40, 1),
    ("float", "Euclidian", 5000, 96, 24, 48, 1),
    ("float", "mips", 4000, 200, 16, 32, 1),
])
def test_build_matches_oracle(dataset, data_type, metric, n, d, R, L, passes):
    ds = dataset(data_type, metric, n=n, d=d, nq=10, R=R, L=L)
    ref, rv, rd = O.vamana_build(ds.base, data_type, metric, R=R, L=L, alpha=1.2, passes=passes, sort_neighbors=False)
    idx, got, (sv, sd) = _build(ds.base, data_type, metric, R, L, passes=passes)
    bad = (valid_graph(got) != valid_graph(ref)).any(axis=1)
    assert not bad.any(), f"{bad.sum()} rows differ, first {np.nonzero(bad)[0][:5]}"
    assert np.array_equal(sv, rv) and np.array_equal(sd, rd)
    assert sum(idx.build_phase_seconds) > 0


def test_build_then_search_roundtrip(dataset):
    # the built graph is searchable in place and reaches high recall (size-independent property)
    ds = dataset("uint8", n=20000, d=128, nq=500, R=32, L=64)
    idx, got, _ = _build(ds.base, "uint8", "Euclidian", 32, 64, stats=False)
    ids, _, _, _ = idx.search(ds.queries, k=10, beam=40)
    gi, gd = O.groundtruth(ds.base, ds.queries, "uint8", k=10)
    from parlayann_b200.harness import recall_with_ties
    assert recall_with_ties(ids, gi, gd, 10) > 0.97


def test_tiny_builds():
    # n below the first batch sizes; max_batch == 0 fix (index.h:208-209)
    for n in (1, 2, 3, 17, 49, 51):
        base = O.gen_points("uint8", n, 32, seed=5)
        ref, _, _ = O.vamana_build(base, "uint8", R=8, L=16, sort_neighbors=False)
        _, got, _ = _build(base, "uint8", "Euclidian", 8, 16, stats=False)
        assert np.array_equal(valid_graph(got), valid_graph(ref)), n


def test_build_errors():
    import parlayann_b200 as P
    base = O.gen_points("uint8", 100, 32)
    idx = P.GraphIndex(base, None, "uint8", "Euclidian", max_degree=8)
    with pytest.raises(ValueError):
        idx.build(R=16, L=16)
    with pytest.raises(P.PannError):
        idx.build(R=8, L=0)


@pytest.mark.parametrize("data_type,metric,n,d,R,L,passes", [
    ("uint8", "Euclidian", 20000, 128, 32, 64, 1), ("uint8", "Euclidian", 6000, 128, 64, 128, 2),
    ("int8", "mips", 6000, 100, 24, 48, 2), ("float", "Euclidian", 5000, 48, 24, 48, 1),
])
def test_build_with_row_cache_matches_oracle(dataset, data_type, metric, n, d, R, L, passes, monkeypatch):
    """K3 with the candidate rows cached in shared memory (the path large builds take) writes the same graph"""
    monkeypatch.setenv("PANN_PRUNE_ROW_CACHE", "1")
    test_build_matches_oracle(dataset, data_type, metric, n, d, R, L, passes)


@pytest.mark.parametrize("row_cache", ["0", "1"])
@pytest.mark.parametrize("data_type,metric,n,d,R,L,passes", [
    ("uint8", "Euclidian", 12000, 128, 32, 64, 2), ("float", "mips", 4000, 200, 16, 32, 1),
])
def test_build_long_list_slab_launch_matches_oracle(dataset, data_type, metric, n, d, R, L, passes, row_cache, monkeypatch):
    """K3's second launch (candidate lists beyond the shared-memory capacity: global slab, bitonic sort, device-side task
    count) is what hubs take in 10M builds; here PANN_PRUNE_CAP forces nearly every list of BOTH modes (visited lists and
    reverse-edge groups) through it, with and without the row cache — same graph and counters as the oracle."""
    monkeypatch.setenv("PANN_PRUNE_CAP", "24")
    monkeypatch.setenv("PANN_PRUNE_ROW_CACHE", row_cache)
    test_build_matches_oracle(dataset, data_type, metric, n, d, R, L, passes)


def test_build_long_list_slab_launch_matches_reference_golden(monkeypatch):
    monkeypatch.setenv("PANN_PRUNE_CAP", "16")
    for name in ("u8_l2", "u8_l2_2pass", "f32_l2"):
        g = golden(name)
        idx, got, (sv, sd) = _build(g.base, g.data_type, g.metric, g.R, g.L, passes=g.passes)
        assert np.array_equal(valid_graph(got), valid_graph(g.z["graph_unsorted"])), name
        assert np.array_equal(sv, g.z["build_visited"]) and np.array_equal(sd, g.z["build_distances"]), name


def test_build_slab_overflow_is_reported(dataset, monkeypatch):
    import parlayann_b200 as P
    ds = dataset("uint8", n=6000, d=128, nq=10, R=32, L=64)
    monkeypatch.setenv("PANN_PRUNE_CAP", "16")
    monkeypatch.setenv("PANN_PRUNE_SLAB_CAP", "64")
    idx = P.GraphIndex(ds.base, None, "uint8", "Euclidian", max_degree=32)
    with pytest.raises(P.PannError, match="PANN_PRUNE_SLAB_CAP"):
        idx.build(R=32, L=64)


def test_build_stats_accounting(dataset):
    """pann_index_build_stats: the device-side B_ins accounting against the oracle's counters: the build's
    distance comparisons = search comparisons + pair distances, so pair distances must equal sum(stats) - search cmps."""
    ds = dataset("uint8", n=8000, d=128, nq=10, R=32, L=64)
    idx, got, (sv, sd) = _build(ds.base, "uint8", "Euclidian", 32, 64)
    st = idx.build_stats()
    assert st["prune_calls"] >= ds.n and st["pair_distances"] > 0
    assert st["algorithmic_bytes"] == st["search_bytes"] + st["prune_bytes"] + ds.n * 2 * 4 * 33
    # every robustPrune call reads at least its candidates' rows: bytes >= calls * (row + graph row)
    assert st["prune_bytes"] >= st["prune_calls"] * (128 + 4 * 33)
    assert st["pair_distances"] < int(sd.astype(np.int64).sum())


@pytest.mark.parametrize("gram", ["1", "2"])
@pytest.mark.parametrize("data_type,metric,n,d,R,L,passes", [
    ("uint8", "Euclidian", 12000, 128, 64, 128, 2), ("int8", "mips", 8000, 128, 32, 64, 1), ("int8", "Euclidian", 8000, 128, 32, 64, 1),
    ("uint8", "mips", 6000, 128, 24, 48, 1),
])
def test_build_with_gram_tile_matches_oracle(dataset, data_type, metric, n, d, R, L, passes, gram, monkeypatch):
    """the candidate-to-candidate distances of robustPrune as tensor-core tiles over the cached rows (PANN_PRUNE_GRAM; measured
    slower than the lazy scan at 10M and off by default): integer distances |xi|^2 + |xj|^2 - 2 xi.xj are the same integers the
    dp4a loop produces, so graphs and counters stay identical — all four 8-bit type / metric combinations"""
    monkeypatch.setenv("PANN_PRUNE_ROW_CACHE", "1")
    monkeypatch.setenv("PANN_PRUNE_GRAM", gram)
    test_build_matches_oracle(dataset, data_type, metric, n, d, R, L, passes)


def test_batch_insert_matches_reference_golden_and_oracle(dataset):
    """pann_vamana_batch_insert = knn_index::batch_insert used directly (vamana/index.h:188-316): subsets into an existing graph,
    given / random order, another max_fraction, re-insertion of inserted vertices — the reference's own graphs (golden), then
    the oracle at a larger size with the graph uploaded in between (pann_index_set_graph)"""
    import os
    import parlayann_b200 as P
    from golden_util import GOLDEN
    from test_oracle_vs_reference import insert_steps
    z = np.load(os.path.join(GOLDEN, "batch_insert.npz"))
    n, d, R, L, seed = (int(x) for x in z["params"])
    base = O.gen_points("uint8", n, d, seed=seed)
    idx = P.GraphIndex(base, None, "uint8", "Euclidian", max_degree=R)
    sv = np.zeros(n, np.uint32); sd = np.zeros(n, np.uint32)
    for i, (ins, alpha, rnd, frac) in enumerate(insert_steps(n)):
        a, b = idx.batch_insert(ins, L=L, alpha=alpha, random_order=rnd, max_fraction=frac, stats=True)
        sv += a; sd += b
        assert np.array_equal(valid_graph(idx.graph()), valid_graph(z[f"g{i}"])), i
    assert np.array_equal(sv, z["bvisited"]) and np.array_equal(sd, z["bdistances"])
    with pytest.raises(P.PannError, match="invalid point"):
        idx.batch_insert(np.array([1, n + 5], np.uint32), L=L)
    idx.batch_insert(np.zeros(0, np.uint32), L=L)   # nothing to insert: the graph stays
    assert np.array_equal(valid_graph(idx.graph()), valid_graph(z["g2"]))
    # larger, int8 MIPS, starting from an oracle-built graph over a prefix
    ds = dataset("int8", "mips", n=9000, d=100, nq=10, R=24, L=48)
    g = np.zeros((ds.n, 25), np.uint32)
    O.vamana_batch_insert(ds.base, g, np.arange(6000, dtype=np.uint32), "int8", "mips", L=48, alpha=1.0, random_order=True)
    idx2 = P.GraphIndex(ds.base, g.copy(), "int8", "mips")
    ins = np.arange(6000, 9000, dtype=np.uint32)
    a, b = idx2.batch_insert(ins, L=48, alpha=1.2, random_order=True, stats=True)
    ra, rb = O.vamana_batch_insert(ds.base, g, ins, "int8", "mips", L=48, alpha=1.2, random_order=True)
    assert np.array_equal(valid_graph(idx2.graph()), valid_graph(g)) and np.array_equal(a, ra) and np.array_equal(b, rb)


def test_tie_rows_sorted_on_the_device_as_std_sort_sorts_them(monkeypatch):
    """the final sort of rows with tied distances: a kernel that runs libstdc++'s introsort per row (vamana_build.cu) against
    the host's std::sort over the same rows (PANN_HOST_SORT=2) and against sorting every row on the host (=1) — 64-entry rows,
    so the quicksort part runs, and 8-bit data of few dimensions, so almost every row has ties"""
    import parlayann_b200 as P
    graphs = {}
    for n, d, R, L in ((30000, 16, 64, 96), (8000, 8, 100, 128)):
        base = O.gen_points("uint8", n, d, seed=9)
        for mode in ("", "2", "1"):
            if mode:
                monkeypatch.setenv("PANN_HOST_SORT", mode)
            else:
                monkeypatch.delenv("PANN_HOST_SORT", raising=False)
            idx = P.GraphIndex(base, None, "uint8", "Euclidian", max_degree=R)
            idx.build(R=R, L=L, alpha=1.2, sort_neighbors=True)
            graphs[mode] = idx.graph()
            idx.close()
        assert (graphs[""][:, 0] > 16).mean() > 0.5, "rows too short to reach the quicksort part"
        assert np.array_equal(graphs[""], graphs["2"]) and np.array_equal(graphs[""], graphs["1"])
