"""The device-side evaluation harness (pann_search_sweep = search_and_parse's loop, check_nn_recall.h:181-268): every row of
the reference's 64-setting sweep — recall (tie-aware, float), average and 99th-percentile visited / distance comparisons —
against the same quantities computed on the host from fetched results, and against the reference's own checkRecall."""
import re

import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


def _ref_stats(c):
    """stats<indexType>::statistics (stats.h:84-92)"""
    s = np.sort(c.astype(np.uint32))
    tot = int(s.astype(np.uint64).sum() & 0xffffffff)
    tot = tot - (1 << 32) if tot >= (1 << 31) else tot
    return int(tot / float(len(s))) & 0xffffffff, int(s[int(np.float64(.99) * np.float64(np.float32(len(s))))])


def test_sweep_rows_match_host_and_reference(dataset, tmp_path):
    import parlayann_b200 as P
    from parlayann_b200 import fileio, harness
    ds = dataset("uint8", n=20000, d=128, nq=500)
    idx = P.GraphIndex(ds.base, ds.graph, "uint8", "Euclidian")
    idx.upload_queries(ds.queries)
    idx.groundtruth_resident(k=100)
    gi, gd = idx.fetch_groundtruth()
    settings = harness.sweep_settings(10, ds.n, ds.R)
    assert len(settings) == 43 + 20 + 1
    rows = harness.search_sweep(idx, settings)
    assert idx.last_timing()[1] >= 3 * 64
    for s, r in zip(settings, rows):
        assert r["status"] == 0
        ids, _, vis, cmps = idx.search(ds.queries, **{k: s[k] for k in ("k", "beam", "cut", "limit", "degree_limit")})
        k = s["k"]
        kk = gi.shape[1]
        truth = (np.arange(kk)[None, :] < k) | (gd == gd[:, k - 1:k])
        correct = int(((ids[:, :, None] == gi[:, None, :].astype(np.uint32)) & truth[:, None, :]).any(axis=1).sum())
        assert r["recall"] == float(np.float32(correct) / np.float32(k * len(ids))), s
        assert (r["avg_cmps"], r["tail_cmps"]) == _ref_stats(cmps) and (r["avg_visited"], r["tail_visited"]) == _ref_stats(vis), s
    # the ground truth can also come from a file (uploaded): same rows
    idx.upload_groundtruth(gi, gd)
    again = harness.search_sweep(idx, settings[:5])
    assert [(r["recall"], r["avg_cmps"], r["tail_visited"]) for r in again] == [(r["recall"], r["avg_cmps"], r["tail_visited"]) for r in rows[:5]]
    # a setting the reference aborts on (k > beam) is reported, the rest of the sweep still runs
    bad = harness.search_sweep(idx, [dict(k=10, beam=5, cut=1.35, limit=ds.n, degree_limit=ds.R), settings[0]])
    assert bad[0]["status"] != 0 and bad[1]["status"] == 0 and bad[1]["recall"] == rows[0]["recall"]
    # the reference's own checkRecall (ref_driver recall) on the same files
    if O.ref_driver_path() is not None:
        files = {k: str(tmp_path / k) for k in ("base", "query", "graph", "gt")}
        fileio.write_bin(files["base"], ds.base); fileio.write_bin(files["query"], ds.queries)
        fileio.write_graph(files["graph"], ds.graph); fileio.write_ibin(files["gt"], gi, gd)
        pick = [0, 7, 20, 42, 43, 50, 62]
        st = ";".join(f"{settings[i]['k']}:{settings[i]['beam']}:{settings[i]['cut']!r}:{settings[i]['limit']}:{settings[i]['degree_limit']}" for i in pick)
        out = O.ref_run(["recall", "-data_type", "uint8", "-dist_func", "Euclidian", "-base", files["base"], "-graph", files["graph"],
                         "-query", files["query"], "-gt", files["gt"], "-settings", st, "-reps", 1])
        ref = re.findall(r"REF_RECALL (\S+) QPS \S+ avg_cmps (\d+) tail_cmps (\d+) avg_visited (\d+) tail_visited (\d+)", out)
        assert len(ref) == len(pick)
        for i, (rec, ac, tc, av, tv) in zip(pick, ref):
            r = rows[i]
            assert abs(float(rec) - r["recall"]) < 1e-6 and (int(ac), int(tc), int(av), int(tv)) == (r["avg_cmps"], r["tail_cmps"], r["avg_visited"], r["tail_visited"]), (i, r)
    rows2, best = harness.search_and_parse(idx, k=10)
    assert len(best) >= 3 and all(b <= r["recall"] for b, r in best)
