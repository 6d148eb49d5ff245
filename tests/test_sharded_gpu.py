"""GPU checks of the multi-GPU layer.  On a 1-GPU box: the world-1 path and the device-pointer merge (two base shards
held by two indexes on the same device, merged by K5 from device memory).  With >= 2 GPUs: a real 2-rank NCCL run."""
import ctypes as C
import os
import socket

import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_world1_paths(dataset):
    import parlayann_b200 as P
    ds = dataset("uint8", "Euclidian", n=20000, d=128, nq=500, R=32, L=64)
    idx = P.GraphIndex(ds.base, ds.graph, "uint8", "Euclidian")
    got = P.sharded.search_sharded(idx, ds.queries, k=10, beam=20)
    ref = O.search(ds.base, ds.graph, ds.queries, "uint8", k=10, beam=20)
    assert np.array_equal(got[0], ref["ids"]) and np.array_equal(got[3], ref["cmps"])
    gi, gd = P.sharded.groundtruth_sharded(idx, ds.queries, k=50)
    oi, od = O.groundtruth(ds.base, ds.queries, "uint8", k=50)
    assert np.array_equal(gd, od)


def test_device_pointer_merge(dataset):
    """what every rank does after the all-gather, emulated on one device"""
    import torch
    import parlayann_b200 as P
    from parlayann_b200 import _lib
    ds = dataset("uint8", "Euclidian", n=20000, d=128, nq=500, R=32, L=64)
    k, S = 100, 3
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    all_i = torch.empty((S, ds.nq, k), dtype=torch.int32, device=dev)
    all_d = torch.empty((S, ds.nq, k), dtype=torch.float32, device=dev)
    keep = []
    for r in range(S):
        lo, hi = P.sharded.shard_bounds(ds.n, r, S)
        idx = P.GraphIndex(ds.base[lo:hi], None, "uint8", "Euclidian", max_degree=4)
        idx.upload_queries(ds.queries)
        idx.groundtruth_resident(k=k, id_offset=lo)
        p_i, p_d, nq, kk = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_uint32()
        _lib.check(L.pann_groundtruth_device_result(idx._h, C.byref(p_i), C.byref(p_d), C.byref(nq), C.byref(kk)))
        assert nq.value == ds.nq and kk.value == k
        all_i[r].copy_(torch.as_tensor(P.sharded._DevView(p_i.value, (ds.nq, k), "<i4"), device=dev))
        all_d[r].copy_(torch.as_tensor(P.sharded._DevView(p_d.value, (ds.nq, k), "<f4"), device=dev))
        keep.append(idx)
    out_i = torch.empty((ds.nq, k), dtype=torch.int32, device=dev)
    out_d = torch.empty((ds.nq, k), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    _lib.check(L.pann_groundtruth_merge_device(0, C.c_uint32(S), C.c_uint64(ds.nq), C.c_uint32(k), C.c_void_p(all_i.data_ptr()),
                                               C.c_void_p(all_d.data_ptr()), C.c_void_p(out_i.data_ptr()), C.c_void_p(out_d.data_ptr()),
                                               C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    whole = P.GraphIndex(ds.base, None, "uint8", "Euclidian", max_degree=4)
    wi, wd = whole.groundtruth(ds.queries, k=k)
    assert np.array_equal(out_d.cpu().numpy(), wd)
    assert np.array_equal(out_i.cpu().numpy(), wi)   # both are sorted by (dist, id): identical including ties


def _nccl_worker(rank, world, port, tmp):
    import torch
    import torch.distributed as dist
    import parlayann_b200 as P
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        n, d, nq, k = 30001, 96, 257, 100
        base = O.gen_points("float_norm", n, d)
        qs = O.gen_points("float_norm", nq, d, stream=1)
        lo, hi = P.sharded.shard_bounds(n, rank, world)
        shard = P.GraphIndex(base[lo:hi], None, "float", "Euclidian", max_degree=4, device=rank)
        gi, gd = (a.copy() for a in P.sharded.groundtruth_sharded(shard, qs, k=k, base_offset=lo))   # views of reused buffers
        oi, od = O.groundtruth(base, qs, "float", k=k)
        assert np.array_equal(gd, od), "sharded ground-truth distances differ from the oracle"
        si, sd = P.sharded.groundtruth_sharded(shard, qs, k=k, base_offset=lo, share_bounds=True, sample_rows=2048)
        assert np.array_equal(sd, gd) and np.array_equal(si, gi), "shared-bound legs changed the sharded ground truth"
        g = O.vamana_build(base[:8000], "float", R=32, L=64, sort_neighbors=False)[0]
        rep = P.GraphIndex(base[:8000], g, "float", "Euclidian", device=rank)
        got = P.sharded.search_sharded(rep, qs, k=10, beam=24)
        ref = O.search(base[:8000], g, qs, "float", k=10, beam=24)
        assert np.array_equal(got[0], ref["ids"]) and np.array_equal(got[2], ref["visited"])
        got0 = P.sharded.search_sharded(rep, qs, k=10, beam=24, root=0)     # answer assembled on rank 0 only
        if rank == 0:
            assert all(np.array_equal(a, b) for a, b in zip(got0, got))
        else:
            assert got0 is None
        open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_nccl_world2(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    mp.spawn(_nccl_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()
