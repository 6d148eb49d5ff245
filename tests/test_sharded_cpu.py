"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: partition bounds, ragged all-gather, sharded search
re-assembly and sharded ground truth (id offsets + merge order).  The compute callbacks are the oracle (checker);
the product's own callbacks are CUDA and are covered by tests/test_sharded_gpu.py."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

import oracle_lib as O


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_shard_bounds():
    from parlayann_b200.sharded import shard_bounds
    for n in (0, 1, 7, 10000, 10000019):
        for w in (1, 2, 3, 4, 8):
            b = [shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        shard_bounds(10, 2, 2)


def _np_merge(list_ids, list_d):
    ids = np.concatenate(list_ids, axis=1); d = np.concatenate(list_d, axis=1)
    k = list_ids[0].shape[1]
    oi = np.empty((ids.shape[0], k), np.int32); od = np.empty((ids.shape[0], k), np.float32)
    for q in range(ids.shape[0]):
        o = np.lexsort((ids[q], d[q]))[:k]
        oi[q], od[q] = ids[q][o], d[q][o]
    return oi, od


class _OracleIndex:
    """stands in for GraphIndex.search in the host-logic test (checker code, not the product path)"""

    def __init__(self, base, graph):
        self.base, self.graph = base, graph

    def search(self, queries, k=10, beam=10, **kw):
        r = O.search(self.base, self.graph, queries, "uint8", k=k, beam=beam)
        return r["ids"], r["dists"], r["visited"], r["cmps"]


def _worker(rank, world, port, tmp):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from parlayann_b200 import sharded as S
        n, d, nq, k = 3001, 32, 101, 10
        base = O.gen_points("uint8", n, d)
        qs = O.gen_points("uint8", nq, d, stream=1)
        # ragged all-gather (uint32 and float rows)
        lo, hi = S.shard_bounds(nq, rank, world)
        counts = [S.shard_bounds(nq, r, world)[1] - S.shard_bounds(nq, r, world)[0] for r in range(world)]
        a = np.arange(nq * 3, dtype=np.uint32).reshape(nq, 3) + 0x80000000
        assert np.array_equal(S.all_gather_rows(a[lo:hi], counts), a)
        # sharded search == unsharded search
        g, _, _ = O.vamana_build(base, "uint8", R=16, L=32, sort_neighbors=False)
        idx = _OracleIndex(base, g)
        got = S.search_sharded(idx, qs, k=k, beam=16)
        ref = idx.search(qs, k=k, beam=16)
        for x, y in zip(got, ref):
            assert np.array_equal(x, y)
        mine = S.search_sharded(idx, qs, gather=False, k=k, beam=16)
        assert np.array_equal(mine[0], ref[0][lo:hi])
        at_root = S.search_sharded(idx, qs, root=1, k=k, beam=16)      # the answer on one rank only
        assert (at_root is None) == (rank != 1)
        if rank == 1:
            assert all(np.array_equal(x, y) for x, y in zip(at_root, ref))
        # sharded ground truth == ground truth of the whole base set (ties: distances identical, ids by (dist,id))
        blo, bhi = S.shard_bounds(n, rank, world)

        def local(queries, kk, off):
            i, dd = O.groundtruth(base[blo:bhi], queries, "uint8", k=kk)
            return (i + off).astype(np.int32), dd
        gi, gd = S.groundtruth_sharded(None, qs, k=20, base_offset=blo, local_fn=local, merge_fn=_np_merge)
        ri, rd = O.groundtruth(base, qs, "uint8", k=20)
        assert np.array_equal(gd, rd)
        for q in range(nq):  # ids equal wherever the distance is unique in the row and not tied with the k-th
            uniq = np.ones(20, bool)
            uniq[1:] &= rd[q][1:] != rd[q][:-1]; uniq[:-1] &= rd[q][:-1] != rd[q][1:]; uniq &= rd[q] != rd[q][-1]
            assert np.array_equal(gi[q][uniq], ri[q][uniq])
        open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_gloo_world2(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
