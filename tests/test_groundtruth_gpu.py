"""K4 brute-force ground truth + K5 shard merge through the C ABI, against the reference goldens and the oracle.
Distances must be bit-identical (integer types exactly; floats because K4 v1 performs the reference's own
operation sequence); ids identical except inside tie groups (the reference's tie order is unspecified,
compute_groundtruth.cpp:68-73)."""
import numpy as np
import pytest

import oracle_lib as O
from golden_util import CASES, golden

pytestmark = pytest.mark.gpu


def _ids_match_up_to_ties(ids, dists, ref_ids):
    for i in range(ids.shape[0]):
        if np.array_equal(ids[i], ref_ids[i]):
            continue
        for dv in np.unique(dists[i]):
            if dv == dists[i, -1]:
                continue
            m = dists[i] == dv
            assert set(ids[i][m].tolist()) == set(ref_ids[i][m].tolist()), i


@pytest.mark.parametrize("name", list(CASES))
def test_groundtruth_matches_reference_golden(name):
    import parlayann_b200 as P
    g = golden(name)
    idx = P.GraphIndex(g.base, None, g.data_type, g.metric, max_degree=4)
    ids, dists = idx.groundtruth(g.queries, k=g.gt_k)
    assert np.array_equal(dists, g.z["gt_dists"]), "ground-truth distances differ from the reference's"
    _ids_match_up_to_ties(ids, dists, g.z["gt_ids"])


@pytest.mark.parametrize("data_type,metric,n,d,nq,k", [
    ("uint8", "Euclidian", 30000, 128, 130, 100),
    ("int8", "mips", 7001, 100, 65, 10),
    ("float", "Euclidian", 9000, 96, 70, 100),
    ("float", "mips", 5000, 200, 33, 1),
    ("uint8", "Euclidian", 5000, 17, 10, 1000),
])
def test_groundtruth_matches_oracle(data_type, metric, n, d, nq, k):
    import parlayann_b200 as P
    kind = "float_norm" if (data_type == "float" and d == 96) else data_type
    base = O.gen_points(kind, n, d, seed=3)
    qs = O.gen_points(kind, nq, d, seed=3, stream=1)
    idx = P.GraphIndex(base, None, data_type, metric, max_degree=4)
    ids, dists = idx.groundtruth(qs, k=k)
    oi, od = O.groundtruth(base, qs, data_type, metric, k=k)
    assert np.array_equal(dists, od)
    # both sort rows by (dist, id); only the cut inside the boundary tie group may differ from the oracle's scan order
    inner = dists < dists[:, -1:]
    assert np.array_equal(ids[inner], oi[inner])
    assert (np.diff(dists, axis=1) >= 0).all()


def test_sharded_groundtruth_merge():
    # base sharded over "GPUs" (here: several indexes on one device), ids offset per shard, K5 merges
    import parlayann_b200 as P
    n, d, nq, k, S = 12000, 64, 50, 20, 4
    base = O.gen_points("uint8", n, d, seed=11)
    qs = O.gen_points("uint8", nq, d, seed=11, stream=1)
    bounds = [0, 2500, 6000, 6100, n]
    parts = []
    for s in range(S):
        idx = P.GraphIndex(base[bounds[s]:bounds[s + 1]], None, "uint8", "Euclidian", max_degree=4)
        parts.append(idx.groundtruth(qs, k=k, id_offset=bounds[s]))
    mi, md = P.groundtruth_merge([p[0] for p in parts], [p[1] for p in parts])
    full = P.GraphIndex(base, None, "uint8", "Euclidian", max_degree=4).groundtruth(qs, k=k)
    assert np.array_equal(mi, full[0]) and np.array_equal(md, full[1])
    oi, od = O.groundtruth(base, qs, "uint8", k=k)
    assert np.array_equal(md, od)


@pytest.mark.parametrize("data_type,kind,metric,n,d,nq,k,sample", [
    ("uint8", "uint8", "Euclidian", 60000, 128, 200, 100, 1024), ("float", "float_norm", "Euclidian", 50000, 96, 130, 100, 2048),
    ("float", "float", "mips", 40000, 200, 100, 10, 512), ("int8", "int8", "Euclidian", 9000, 128, 64, 20, 4096),
])
def test_sharded_groundtruth_shared_bounds(data_type, kind, metric, n, d, nq, k, sample):
    """the two-leg sharded job (sharded.groundtruth_sharded's GPU path, here with the shards on one device): every shard
    takes a sample, the merged sample's k-th distance bounds every shard's filter for the rest of its rows; the merged
    result must be the unsharded one bit for bit, although the bounded shards keep fewer than k rows each"""
    import torch
    import parlayann_b200 as P
    base = O.gen_points(kind, n, d, seed=5)
    qs = O.gen_points(kind, nq, d, seed=5, stream=1)
    bounds = [0, n // 2, n // 2 + 300, n]          # the middle shard is smaller than the sample
    shards = [P.GraphIndex(base[bounds[s]:bounds[s + 1]], None, data_type, metric, max_degree=4) for s in range(3)]
    srows = []
    samples = []
    for s, idx in enumerate(shards):
        idx.upload_queries(qs)
        r = min(idx.n, sample)
        idx.groundtruth_range(k=k, id_offset=bounds[s], row_lo=0, row_hi=r)
        srows.append(r)
        samples.append(idx.fetch_groundtruth())
    mi, md = P.groundtruth_merge([p[0] for p in samples], [p[1] for p in samples])
    assert (mi >= 0).all()
    cap = torch.from_numpy(np.ascontiguousarray(md[:, k - 1])).cuda()
    parts, short = [], 0
    for s, idx in enumerate(shards):
        if srows[s] < idx.n:
            idx.groundtruth_range(k=k, id_offset=bounds[s], row_lo=srows[s], row_hi=0, keep_state=True, d_tau_cap=cap.data_ptr(), seen_rows=sum(srows))
            if idx.n - srows[s] >= 1024:
                assert idx.groundtruth_stats()[0] > 0, "the bounded leg must run on the tensor cores"
        parts.append(idx.fetch_groundtruth())
    gi, gd = P.groundtruth_merge([p[0] for p in parts], [p[1] for p in parts])
    full = P.GraphIndex(base, None, data_type, metric, max_degree=4).groundtruth(qs, k=k)
    assert np.array_equal(gd.view(np.uint32), full[1].view(np.uint32)) and np.array_equal(gi, full[0])
    # a bounded shard does not chase rows that cannot be in the global top-k: its own k-th is not tighter than the bound
    unb = shards[0].groundtruth(qs, k=k, id_offset=0)
    assert (np.where(parts[0][0] >= 0, parts[0][1], np.inf) >= unb[1]).all()
    with pytest.raises(P.PannError):
        shards[0].groundtruth_range(k=k, row_lo=100)       # not a tile boundary
    with pytest.raises(P.PannError):
        shards[0].groundtruth_range(k=k + 1, row_lo=256, keep_state=True)   # no kept state of that shape


def test_groundtruth_recall_of_itself_is_one():
    import parlayann_b200 as P
    from parlayann_b200.harness import recall_with_ties
    base = O.gen_points("uint8", 4000, 32, seed=2)
    qs = O.gen_points("uint8", 40, 32, seed=2, stream=1)
    ids, dists = P.GraphIndex(base, None, "uint8", "Euclidian", max_degree=4).groundtruth(qs, k=30)
    oi, od = O.groundtruth(base, qs, "uint8", k=30)
    assert recall_with_ties(ids[:, :10], oi, od, 10) == 1.0


def test_groundtruth_errors():
    import parlayann_b200 as P
    base = O.gen_points("uint8", 50, 32)
    idx = P.GraphIndex(base, None, "uint8", "Euclidian", max_degree=4)
    with pytest.raises(P.PannError):
        idx.groundtruth(base[:5], k=51)
    with pytest.raises(P.PannError):
        idx.groundtruth(base[:5], k=0)
    ids, dists = idx.groundtruth(base[:5], k=50)
    assert (ids[:, 0] == np.arange(5)).all() and (dists[:, 0] == 0).all()


# ---- K4 v2 (tcgen05 / TMA GEMM with the fused threshold filter) against K4 v1 (CUDA cores, PANN_GT_V1=1)
def _both_paths(idx, queries, k, monkeypatch):
    monkeypatch.setenv("PANN_GT_V1", "0")
    a = idx.groundtruth(queries, k=k)
    n_mma = idx.groundtruth_stats()[0]
    monkeypatch.setenv("PANN_GT_V1", "1")
    b = idx.groundtruth(queries, k=k)
    assert idx.groundtruth_stats()[0] == 0, "PANN_GT_V1=1 must stay on the CUDA cores"
    return a, b, n_mma


@pytest.mark.parametrize("data_type,metric,kind,n,d,nq,k", [
    ("uint8", "Euclidian", "uint8", 30000, 128, 300, 100), ("uint8", "mips", "uint8", 20000, 128, 130, 10),
    ("int8", "Euclidian", "int8", 20000, 128, 129, 50), ("int8", "mips", "int8", 20011, 160, 100, 100),
    ("float", "Euclidian", "float_norm", 40000, 96, 300, 100), ("float", "mips", "float", 20000, 200, 130, 20),
    ("float", "Euclidian", "float", 20000, 100, 257, 1), ("uint8", "Euclidian", "uint8", 20000, 200, 64, 1000),
])
def test_mma_path_identical_to_cuda_core_path(data_type, metric, kind, n, d, nq, k, monkeypatch):
    import parlayann_b200 as P
    base = O.gen_points(kind, n, d)
    qs = O.gen_points(kind, nq, d, stream=1)
    idx = P.GraphIndex(base, None, data_type, metric, max_degree=4)
    (gi, gd), (vi, vd), n_mma = _both_paths(idx, qs, k, monkeypatch)
    assert n_mma > 0, "the tcgen05 kernel did not run for this shape"
    assert np.array_equal(gd.view(np.uint32), vd.view(np.uint32))   # bit-identical floats
    assert np.array_equal(gi, vi)                                   # both sorted by (dist, id): same ids, ties included
    oi, od = O.groundtruth(base, qs, data_type, metric, k=k)
    assert np.array_equal(gd, od)


def test_mma_path_overflow_falls_back(monkeypatch):
    """rows ordered so that every row is closer than all rows before it: the thresholds stay loose, the candidate
    lists overflow, and the chunk is taken again on the CUDA cores — the result must not change"""
    import parlayann_b200 as P
    n, d, nq, k = 30000, 128, 64, 10
    lvl = (255 - (np.arange(n) * 200 // n)).astype(np.uint8)            # 255 ... 56, non-increasing
    base = np.repeat(lvl[:, None], d, axis=1)
    base[:, 0] = (np.arange(n) % 7).astype(np.uint8)                    # break exact ties a little
    qs = np.zeros((nq, d), np.uint8)
    qs[:, 1] = np.arange(nq, dtype=np.uint8)
    idx = P.GraphIndex(base, None, "uint8", "Euclidian", max_degree=4)
    (gi, gd), (vi, vd), _ = _both_paths(idx, qs, k, monkeypatch)
    assert np.array_equal(gd, vd) and np.array_equal(gi, vi)
    oi, od = O.groundtruth(base, qs, "uint8", "Euclidian", k=k)
    assert np.array_equal(gd, od)
