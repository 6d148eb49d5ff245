"""where the wall time of groundtruth_sharded goes (C3): local top-k, all-gathers, merge, copy back — timed with a device
synchronisation after each phase, max over ranks"""
import ctypes as C, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import parlayann_b200 as P
from parlayann_b200 import _lib
from parlayann_b200._lib import check
from parlayann_b200.sharded import shard_bounds, _DevView

n = int(os.environ.get("GT_N", 10_000_000)); d = 96; nq = 10_000; k = 100
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
lo, hi = shard_bounds(n, rank, world)
idx = P.GraphIndex(data_type="float", metric="Euclidian", max_degree=4, device=lr, synthetic=dict(n=hi - lo, d=d, kind="float_norm", seed=1234, stream=0, row0=lo))
idx.synthetic_queries(nq, kind="float_norm", seed=1234, stream=1)
L = _lib.lib()
dev = torch.device("cuda", lr)
for it in range(4):
    ph = {}
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t = time.perf_counter()
    idx.groundtruth_resident(k=k, id_offset=lo)
    ph["local_wall"] = time.perf_counter() - t; ph["local_device"] = idx.last_timing()[0] * 1e-3
    t = time.perf_counter()
    p_ids, p_d, nqv, kk = C.c_void_p(), C.c_void_p(), C.c_uint64(), C.c_uint32()
    check(L.pann_groundtruth_device_result(idx._h, C.byref(p_ids), C.byref(p_d), C.byref(nqv), C.byref(kk)))
    loc_i = torch.as_tensor(_DevView(p_ids.value, (nq, k), "<i4"), device=dev)
    loc_d = torch.as_tensor(_DevView(p_d.value, (nq, k), "<f4"), device=dev)
    all_i = torch.empty((world, nq, k), dtype=torch.int32, device=dev); all_d = torch.empty((world, nq, k), dtype=torch.float32, device=dev)
    torch.cuda.synchronize(); ph["wrap_alloc"] = time.perf_counter() - t; t = time.perf_counter()
    dist.all_gather_into_tensor(all_i, loc_i); dist.all_gather_into_tensor(all_d, loc_d)
    torch.cuda.synchronize(); ph["all_gather"] = time.perf_counter() - t; t = time.perf_counter()
    out_i = torch.empty((nq, k), dtype=torch.int32, device=dev); out_d = torch.empty((nq, k), dtype=torch.float32, device=dev)
    check(L.pann_groundtruth_merge_device(dev.index, C.c_uint32(world), C.c_uint64(nq), C.c_uint32(k), C.c_void_p(all_i.data_ptr()), C.c_void_p(all_d.data_ptr()),
                                          C.c_void_p(out_i.data_ptr()), C.c_void_p(out_d.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize(); ph["merge"] = time.perf_counter() - t; t = time.perf_counter()
    a, b = out_i.cpu().numpy(), out_d.cpu().numpy()
    ph["copy_back"] = time.perf_counter() - t
    v = torch.tensor(list(ph.values()), dtype=torch.float64, device=dev)
    dist.all_reduce(v, op=dist.ReduceOp.MAX)
    if rank == 0 and it:
        print(json.dumps({"gpus": world, **{kk_: round(float(x) * 1e3, 3) for kk_, x in zip(ph.keys(), v.tolist())}}), flush=True)
dist.barrier(); dist.destroy_process_group()
