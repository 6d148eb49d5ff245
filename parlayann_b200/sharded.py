"""Multi-GPU layer: one process per GPU (torch.distributed), only where the path shards.

* search       — queries are independent (beamSearch.h:556): the index is replicated, the query batch is cut into
                 `world` contiguous slices, every rank searches its slice; no data-path collective.  The slices are
                 put back together with one all-gather of nq*k ids (400 KB at 10 000 x 10) when the caller wants the
                 whole answer on every rank.
* ground truth — the BASE rows are sharded (compute_groundtruth.cpp:22-60 scans all of them per query): every rank
                 computes the top-k of all queries against its shard with global ids (id_offset), the ranks exchange
                 the [nq, k] blocks with one all-gather (NCCL over NVLink on GPUs) and each merges the `world` sorted
                 rows per query with K5 under the (dist, id) order.
* Vamana build — does not shard (each batch reads the graph the previous batch wrote): one GPU, replicas only.

The partition helpers are pure functions (tested on CPU with world_size-2 gloo groups); the compute callbacks default
to the CUDA library and can be replaced by a checker in tests.
"""
import ctypes as C
import os

import numpy as np

from . import _lib
from ._lib import check


def shard_bounds(n, rank, world):
    """Contiguous slice [lo, hi) of n items for `rank`: the first n % world ranks get one more."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    q, r = divmod(int(n), int(world))
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def _dist():
    import torch.distributed as dist
    return dist


def _world(group=None):
    dist = _dist()
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def _comm_device(group=None):
    import torch
    dist = _dist()
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")


def all_gather_rows(local, counts, group=None):
    """Concatenation over ranks of `local` (rows of rank r: counts[r]); ragged slices are padded for the collective."""
    import torch
    dist = _dist()
    rank, world = _world(group)
    if world == 1:
        return local
    dev = _comm_device(group)
    m = max(counts)
    pad = np.zeros((m,) + local.shape[1:], dtype=local.dtype)
    pad[:local.shape[0]] = local
    view = pad.view(np.int32) if pad.dtype == np.uint32 else pad   # torch has no uint32 collectives
    t = torch.from_numpy(view).to(dev)
    outs = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(outs, t, group=group)
    parts = [o.cpu().numpy().view(local.dtype)[:counts[r]] for r, o in enumerate(outs)]
    return np.concatenate(parts, axis=0)


def search_sharded(index, queries, gather=True, group=None, root=None, **qp):
    """qsearchAll over a query batch cut across the ranks of `group`.  `index` is this rank's replica
    (GraphIndex).  Returns (ids, dists, visited, cmps) for all queries when gather=True, else this rank's slice.

    With an NCCL group the slice's results never leave the device before the exchange: queries H2D, search, one
    all-gather per result array straight from the library's device buffers (pann_search_device_result), one D2H of the
    assembled answer into pinned buffers kept on the index.  root=r: only rank r assembles (and returns) the answer — what
    a single driving process needs; the other ranks return None and skip the device-to-host copy of the whole batch.
    The gathered arrays live in page-locked buffers kept on the index and are overwritten by the next call: copy what must
    outlive it."""
    rank, world = _world(group)
    nq = queries.shape[0]
    lo, hi = shard_bounds(nq, rank, world)
    k = int(qp.get("k", 10))
    if gather and world > 1 and _dist().get_backend(group) == "nccl":
        return _search_sharded_nccl(index, queries, lo, hi, k, group, qp, root)
    if world == 1 and nq:   # one rank: the whole batch, into page-locked result buffers kept on the index (as the NCCL path does)
        from .index import pinned_empty
        cache = index.__dict__.setdefault("_sharded_out", {})   # per batch shape: freeing page-locked memory costs milliseconds
        out = cache.get((nq, k))
        if out is None:
            if len(cache) >= 4:
                cache.clear()
            out = (pinned_empty((nq, k), np.uint32), pinned_empty((nq, k), np.float32), pinned_empty((nq,), np.uint32), pinned_empty((nq,), np.uint32))
            cache[(nq, k)] = out
        return index.search(queries, out=out, **qp)
    if hi > lo:
        ids, dists, vis, cmps = index.search(queries[lo:hi], **qp)
    else:  # more ranks than queries
        ids, dists = np.empty((0, k), np.uint32), np.empty((0, k), np.float32)
        vis, cmps = np.empty(0, np.uint32), np.empty(0, np.uint32)
    if not gather or world == 1:
        return ids, dists, vis, cmps
    counts = [shard_bounds(nq, r, world)[1] - shard_bounds(nq, r, world)[0] for r in range(world)]
    out = tuple(all_gather_rows(a, counts, group) for a in (ids, dists, vis, cmps))
    return out if root is None or rank == root else None      # host-logic path (gloo): same contract as the NCCL gather to one rank


def _search_sharded_nccl(index, queries, lo, hi, k, group, qp, root=None):
    import torch
    dist = _dist()
    rank, world = _world(group)
    nq = queries.shape[0]
    counts = [shard_bounds(nq, r, world)[1] - shard_bounds(nq, r, world)[0] for r in range(world)]
    m = max(counts)
    dev = torch.device("cuda", torch.cuda.current_device())
    cache = getattr(index, "_shard_bufs", None)
    if cache is None or cache["key"] != (world, m, k):
        cache = {"key": (world, m, k),
                 "g": [torch.empty((world, m, k), dtype=torch.int32, device=dev), torch.empty((world, m, k), dtype=torch.float32, device=dev),
                       torch.empty((world, m), dtype=torch.int32, device=dev), torch.empty((world, m), dtype=torch.int32, device=dev)],
                 "l": [torch.zeros((m, k), dtype=torch.int32, device=dev), torch.zeros((m, k), dtype=torch.float32, device=dev),
                       torch.zeros((m,), dtype=torch.int32, device=dev), torch.zeros((m,), dtype=torch.int32, device=dev)]}
        cache["h"] = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in cache["g"]]
        index._shard_bufs = cache
    n_loc = hi - lo
    if n_loc > 0:
        index.upload_queries(queries[lo:hi])
        index.search_resident(**qp)
        L = _lib.lib()
        p = [C.c_void_p() for _ in range(4)]
        rn, rk = C.c_uint64(), C.c_uint32()
        check(L.pann_search_device_result(index._h, C.byref(p[0]), C.byref(p[1]), C.byref(p[2]), C.byref(p[3]), C.byref(rn), C.byref(rk)))
        views = [torch.as_tensor(_DevView(p[0].value, (n_loc, k), "<i4"), device=dev), torch.as_tensor(_DevView(p[1].value, (n_loc, k), "<f4"), device=dev),
                 torch.as_tensor(_DevView(p[2].value, (n_loc,), "<i4"), device=dev), torch.as_tensor(_DevView(p[3].value, (n_loc,), "<i4"), device=dev)]
        for dst, src in zip(cache["l"], views):
            dst[:n_loc].copy_(src)          # equal slices: the padded staging buffer IS the slice (one D2D copy, NVLink-sized)
    if root is None:
        for g, l in zip(cache["g"], cache["l"]):
            dist.all_gather_into_tensor(g, l, group=group)
    else:
        for g, l in zip(cache["g"], cache["l"]):
            dist.gather(l, [g[r] for r in range(world)] if rank == root else None, dst=dist.get_global_rank(group, root) if group is not None else root, group=group)
        if rank != root:
            torch.cuda.current_stream().synchronize()
            return None
    for h, g in zip(cache["h"], cache["g"]):
        h.copy_(g, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    outs = []
    for h, dt in zip(cache["h"], (np.uint32, np.float32, np.uint32, np.uint32)):
        a = h.numpy()
        if all(c == m for c in counts):
            a = a.reshape((world * m,) + a.shape[2:])
        else:
            a = np.concatenate([a[r, :counts[r]] for r in range(world)], axis=0)
        outs.append(a.view(dt) if a.dtype != dt else a)
    return tuple(outs)


class _DevView:
    """a device pointer as a __cuda_array_interface__ object (read by torch.as_tensor without a copy)"""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def _gt_gather_merge(L, index, dist, group, world, k, dev, torch):
    """all-gather of this rank's device-resident [nq,k] ids / distances + K5 merge on the device -> (ids, dists) tensors"""
    p_ids, p_d, nq, kk = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_uint32()
    check(L.pann_groundtruth_device_result(index._h, C.byref(p_ids), C.byref(p_d), C.byref(nq), C.byref(kk)))
    nq = nq.value
    loc_i = torch.as_tensor(_DevView(p_ids.value, (nq, k), "<i4"), device=dev)
    loc_d = torch.as_tensor(_DevView(p_d.value, (nq, k), "<f4"), device=dev)
    all_i = torch.empty((world, nq, k), dtype=torch.int32, device=dev)
    all_d = torch.empty((world, nq, k), dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(all_i, loc_i, group=group)
    dist.all_gather_into_tensor(all_d, loc_d, group=group)
    out_i = torch.empty((nq, k), dtype=torch.int32, device=dev)
    out_d = torch.empty((nq, k), dtype=torch.float32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    check(L.pann_groundtruth_merge_device(dev.index, C.c_uint32(world), C.c_uint64(nq), C.c_uint32(k), C.c_void_p(all_i.data_ptr()),
                                          C.c_void_p(all_d.data_ptr()), C.c_void_p(out_i.data_ptr()), C.c_void_p(out_d.data_ptr()),
                                          C.c_void_p(st)))
    return out_i, out_d


GT_SAMPLE_ROWS = 4096   # rows of every shard that make the shared first bound (a multiple of the 256-row MMA tile)


def groundtruth_sharded(index, queries=None, k=100, base_offset=0, group=None, local_fn=None, merge_fn=None, share_bounds=None,
                        sample_rows=GT_SAMPLE_ROWS):
    """compute_groundtruth with the base rows sharded over the ranks: `index` holds this rank's base shard whose
    first row has global id `base_offset`.  Every rank returns the global (ids [nq,k] int32, dists [nq,k] float32) — on the
    GPU path as views of page-locked buffers kept on the index (copy them if they must outlive the next call).

    Default path (GPU): local top-k stays on the device, NCCL all-gather of the [nq,k] id and distance blocks,
    K5 merge on the device, one D2H copy of the result.  With share_bounds (or PANN_GT_SHARE=1; off by default: measured
    on 10 000 x 10M x 96 f32 it saves 0.6 ms of K4 per shard at 8 GPUs and costs 0.8 ms of extra exchange) every shard first takes a sample_rows-row
    sample, the samples are merged the same way, and the k-th distance of the merged sample bounds every shard's filter
    for the rest of its rows (a shard on its own only knows the k-th distance of ITS rows, which is ~world times looser
    in rank): same result, fewer survivors to re-rank per shard.
    local_fn(queries, k, base_offset) -> (ids, dists) and merge_fn(list_ids, list_dists) -> (ids, dists) replace the
    compute for host-logic tests."""
    rank, world = _world(group)
    if local_fn is not None or merge_fn is not None:
        ids, dists = local_fn(queries, k, base_offset)
        if world == 1:
            return merge_fn([ids], [dists])
        nq = ids.shape[0]
        gi = all_gather_rows(ids, [nq] * world, group).reshape(world, nq, k)
        gd = all_gather_rows(dists, [nq] * world, group).reshape(world, nq, k)
        return merge_fn(list(gi), list(gd))
    import torch
    L = _lib.lib()
    if queries is not None:
        index.upload_queries(queries)
    if world == 1:
        index.groundtruth_resident(k=k, id_offset=base_offset)
        return index.fetch_groundtruth()
    dist = _dist()
    if dist.get_backend(group) != "nccl":
        raise RuntimeError("groundtruth_sharded: the GPU path needs an NCCL process group")
    dev = torch.device("cuda", torch.cuda.current_device())
    if share_bounds is None:
        share_bounds = os.environ.get("PANN_GT_SHARE", "0") == "1"
    legs = []
    if share_bounds:
        n_loc = index.n
        s_rows = min(n_loc, max(256, sample_rows // 256 * 256))
        if s_rows < n_loc:
            s_rows = s_rows // 256 * 256
        index.groundtruth_range(k=k, id_offset=base_offset, row_lo=0, row_hi=s_rows)
        legs.append(index.last_timing()[0])
        seen = torch.tensor([s_rows], dtype=torch.int64, device=dev)
        dist.all_reduce(seen, group=group)
        _, samp_d = _gt_gather_merge(L, index, dist, group, world, k, dev, torch)
        cap = samp_d[:, k - 1].contiguous()          # +inf where the merged sample holds fewer than k rows: no bound
        torch.cuda.current_stream().synchronize()    # the library runs on its own stream
        if s_rows < n_loc:
            index.groundtruth_range(k=k, id_offset=base_offset, row_lo=s_rows, row_hi=0, keep_state=True, d_tau_cap=cap.data_ptr(),
                                    seen_rows=int(seen.item()))
            legs.append(index.last_timing()[0])
    else:
        index.groundtruth_resident(k=k, id_offset=base_offset)
        legs.append(index.last_timing()[0])
    index._gt_leg_ms = legs            # device time of this rank's K4 legs (diagnostics, bench.py)
    out_i, out_d = _gt_gather_merge(L, index, dist, group, world, k, dev, torch)
    nq = out_i.shape[0]
    # results come back through pinned buffers kept on the index (a pageable .cpu() of 8 MB takes 5 ms, this 0.3 ms)
    pin = getattr(index, "_gt_pinned", None)
    if pin is None or pin[0].shape != (nq, k):
        pin = (torch.empty((nq, k), dtype=torch.int32).pin_memory(), torch.empty((nq, k), dtype=torch.float32).pin_memory())
        index._gt_pinned = pin
    pin[0].copy_(out_i, non_blocking=True)
    pin[1].copy_(out_d, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return pin[0].numpy(), pin[1].numpy()      # views of the page-locked buffers: overwritten by the next call on this index
