"""Evaluation harness: the reference's tie-aware recall and its search sweep.

Reference: algorithms/utils/check_nn_recall.h:17-125 (checkRecall), :181-268 (search_and_parse),
algorithms/utils/parse_results.h:143-148, :193-218 (result rows, recall buckets), algorithms/utils/stats.h:84-92.
The sweep itself runs on the device (pann_search_sweep): search, recall and counter statistics of every setting stay
there and one copy brings all rows back."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import QueryParams, SweepResult, check

BEAMS = [10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 22, 24, 26, 28, 30, 32, 34, 36, 38, 40, 45, 50, 55, 60, 65, 70, 80, 90, 100,
         120, 140, 160, 180, 200, 225, 250, 275, 300, 375, 500, 750, 1000]
LIMITS = [10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 28, 30, 35]
BUCKETS = [.1, .2, .3, .4, .5, .6, .7, .75, .8, .85, .9, .93, .95, .97, .98, .99, .995, .999, .9995, .9999, .99995, .99999]


def recall_with_ties(neighbors, gt_ids, gt_dists, k):
    """check_nn_recall.h:83-109 on the host: the ground-truth set of a query is its first k ids plus every later id
    whose distance equals the k-th distance; recall = (# reported ids in that set) / (k * nq)."""
    neighbors = np.asarray(neighbors)
    nq = gt_ids.shape[0]
    correct = 0
    for i in range(nq):
        last = gt_dists[i, k - 1]
        extra = np.nonzero(gt_dists[i, k:] == last)[0]
        truth = set(gt_ids[i, :k].tolist())
        truth.update(gt_ids[i, k + extra].tolist())
        correct += len(truth.intersection(set(neighbors[i, :k].tolist())))
    return float(correct) / float(k * nq)


def counter_stats(c):
    """stats<indexType>::statistics (stats.h:84-92): (int)sum / n and the 0.99 quantile element."""
    s = np.sort(np.asarray(c, dtype=np.int64))
    return int(s.sum() // len(s)), int(s[int(0.99 * len(s))])


def sweep_settings(k, n, max_degree, fixed_beam_width=0):
    """The QueryParams of search_and_parse (check_nn_recall.h:214-257) for one k, in its order: 43 beam widths at cut 1.35
    (passed through a float, :231), 20 visit limits with beam max(l, k) and degree limit min(maxDeg, 5 l), and the
    'best accuracy' setting (k=100, beam 1000, cut 10).  fixed_beam_width (-Q): that one setting, five times (:221-226)."""
    if fixed_beam_width:
        return [dict(k=k, beam=fixed_beam_width, cut=1.35, limit=n, degree_limit=max_degree)] * 5
    cut_f = float(np.float32(1.35))
    out = [dict(k=k, beam=b, cut=cut_f, limit=n, degree_limit=max_degree) for b in BEAMS if b >= k]
    out += [dict(k=k, beam=max(l, k), cut=1.35, limit=l, degree_limit=min(max_degree, 5 * l)) for l in LIMITS]
    out.append(dict(k=100, beam=1000, cut=10.0, limit=n, degree_limit=max_degree))
    return out


def search_sweep(index, settings, starts=None):
    """pann_search_sweep over the index's RESIDENT queries and ground truth: one dict per setting with the fields of the
    reference's nn_result (parse_results.h:104-131); rows the reference would abort on carry status != 0."""
    st = np.asarray([index.start_point] if starts is None else starts, dtype=np.uint32)
    n = len(settings)
    qps = (QueryParams * n)(*[QueryParams(int(s["k"]), int(s["beam"]), float(s["cut"]), int(s["limit"]), int(s["degree_limit"])) for s in settings])
    res = (SweepResult * n)()
    check(_lib.lib().pann_search_sweep(index._h, st.ctypes.data_as(C.c_void_p), C.c_uint32(len(st)), qps, C.c_uint32(n), res))
    rows = []
    for s, r in zip(settings, res):
        qps_ = index.nq / (r.device_ms * 1e-3) if r.device_ms > 0 else 0.0
        rows.append(dict(s, recall=float(np.float32(r.recall)), avg_cmps=r.avg_cmps, tail_cmps=r.tail_cmps, avg_visited=r.avg_visited,
                         tail_visited=r.tail_visited, device_ms=r.device_ms, QPS=qps_, status=r.status))
    return rows


def parse_result(rows, buckets=BUCKETS):
    """parse_results.h:193-218: per recall bucket [b, next b] the row with the highest QPS"""
    out = []
    ok = [r for r in rows if r["status"] == 0]
    for i, b in enumerate(buckets):
        cand = [r for r in ok if r["recall"] >= b]
        if i != len(buckets) - 1 and cand:
            cand = [r for r in cand if r["recall"] <= buckets[i + 1]]
        if cand:
            out.append((b, max(cand, key=lambda r: r["QPS"])))
    return out


def search_and_parse(index, k=10, fixed_beam_width=0, verbose=False):
    """search_and_parse (check_nn_recall.h:181-268) for an index with resident queries and ground truth; prints the
    reference's lines and returns (all rows, bucket rows)"""
    rows = search_sweep(index, sweep_settings(k, index.n, index.max_degree, fixed_beam_width))
    if verbose:
        for r in rows:
            if r["status"] == 0:
                print(f"search: Q={r['beam']}, k={r['k']}, limit={r['limit']}, recall={r['recall']:.6g}, visited={r['avg_visited']}, "
                      f"comparisons={r['avg_cmps']}, QPS={r['QPS']:.4g}")
    best = parse_result(rows)
    for b, r in best:
        print(f"For {r['k']}@{r['k']} recall = {r['recall']:.6g}, QPS = {r['QPS']:.6g}, Q = {r['beam']}, cut = {r['cut']:.6g}, visited limit = {r['limit']}, "
              f"degree limit: {r['degree_limit']}, average visited = {r['avg_visited']}, average cmps = {r['avg_cmps']}")
    return rows, best
