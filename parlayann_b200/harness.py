"""Host-side evaluation harness: the reference's tie-aware recall and its search sweep.

Reference: algorithms/utils/check_nn_recall.h:17-125 (checkRecall), :181-268 (search_and_parse),
algorithms/utils/stats.h:84-92 (avg / p99 of the per-query counters)."""
import numpy as np


def recall_with_ties(neighbors, gt_ids, gt_dists, k):
    """check_nn_recall.h:83-109: the ground-truth set of a query is its first k ids plus every later id
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


def sweep_settings(k, n, max_degree):
    """The QueryParams settings of search_and_parse (check_nn_recall.h:214-257) for one k."""
    beams = [i for i in range(10, 41) if i >= k] + [b for b in (45, 50, 55, 60, 65, 70, 80, 90, 100, 120, 140, 160, 180,
                                                           200, 225, 250, 275, 300, 375, 500, 750, 1000) if b >= k]
    out = []
    for b in beams:
        out.append(dict(k=k, beam=b, cut=1.35, limit=n, degree_limit=max_degree))
    return out
