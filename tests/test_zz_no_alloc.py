"""pann_groundtruth_static (include/pann.h): the ground-truth entry point that allocates nothing — all working memory is a
file-scope device array inside libpann_cuda.so.  It must return what pann_groundtruth returns, bit for bit, on
successive calls (the array keeps the previous call's contents), and refuse sizes beyond the array.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


def _static_gt(P, L, torch, idx, qs_dev, nq, row_bytes, k, id_offset=0):
    out_i = torch.full((nq, k), -7, dtype=torch.int32, device="cuda")
    out_d = torch.full((nq, k), -7.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    before = (torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0])
    rc = L.pann_groundtruth_static(idx._h, C.c_void_p(qs_dev.data_ptr()), C.c_uint64(nq), C.c_uint64(row_bytes), C.c_uint32(k),
                                   C.c_int64(id_offset), C.c_void_p(out_i.data_ptr()), C.c_void_p(out_d.data_ptr()), C.c_void_p(0))
    if rc != 0:
        raise P.PannError(L.pann_last_error().decode())
    torch.cuda.synchronize()
    after = (torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0])
    _static_gt.mem.append((before, after))
    return out_i.cpu().numpy(), out_d.cpu().numpy()


_static_gt.mem = []


@pytest.mark.parametrize("data_type,kind,metric,n,d,nq,k", [
    ("uint8", "uint8", "Euclidian", 20000, 128, 300, 100), ("float", "float", "mips", 9000, 100, 129, 10),
    ("int8", "int8", "Euclidian", 5000, 50, 64, 20),
])
def test_static_scratch_groundtruth_equals_allocating_entry_point(data_type, kind, metric, n, d, nq, k):
    import torch
    import parlayann_b200 as P
    from parlayann_b200 import _lib
    L = _lib.lib()
    base = O.gen_points(kind, n, d, seed=21)
    idx = P.GraphIndex(base, None, data_type, metric, max_degree=4)
    for call in range(2):                       # two calls in a row: the second one finds the first one's scratch contents
        qs = O.gen_points(kind, nq - call * 5, d, seed=21, stream=1 + call)
        ref_i, ref_d = idx.groundtruth(qs, k=k, id_offset=call * 1000)      # the existing entry point
        qd = torch.from_numpy(qs).cuda()
        gi, gd = _static_gt(P, L, torch, idx, qd, qs.shape[0], qs.strides[0], k, id_offset=call * 1000)
        assert np.array_equal(gd.view(np.uint32), ref_d.view(np.uint32)), f"call {call}: distances differ"
        assert np.array_equal(gi, ref_i), f"call {call}: ids differ"
    # no device memory taken by the second call (the first may load the kernels' code): neither from torch nor the driver
    before, after = _static_gt.mem[-1]
    assert after[0] == before[0] and after[1] >= before[1], (before, after)
    # strided device rows (the leading columns of a wider matrix)
    qs0 = O.gen_points(kind, nq, d, seed=21, stream=3)
    wide = torch.zeros((nq, d + 24), dtype=qd.dtype, device="cuda")
    wide[:, :d] = torch.from_numpy(qs0).cuda()
    ref_i, ref_d = idx.groundtruth(qs0, k=k)
    gi, gd = _static_gt(P, L, torch, idx, wide, nq, wide.stride(0) * wide.element_size(), k)
    assert np.array_equal(gd.view(np.uint32), ref_d.view(np.uint32)) and np.array_equal(gi, ref_i)


def test_static_scratch_limit_is_enforced():
    import torch
    import parlayann_b200 as P
    from parlayann_b200 import _lib
    L = _lib.lib()
    lim = C.c_uint64(0)
    assert L.pann_groundtruth_static_limit(C.byref(lim)) == 0 and lim.value == 16 << 20
    base = O.gen_points("uint8", 3000, 128, seed=3)
    idx = P.GraphIndex(base, None, "uint8", "Euclidian", max_degree=4)
    nq, k = 20000, 100                          # 20000 * (800 + 128 + 256) bytes > 16 MiB
    qd = torch.zeros((nq, 128), dtype=torch.uint8, device="cuda")
    with pytest.raises(P.PannError, match="static array holds"):
        _static_gt(P, L, torch, idx, qd, nq, 128, k)
    with pytest.raises(P.PannError):
        _static_gt(P, L, torch, idx, qd, 10, 64, k)          # stride smaller than a row
    with pytest.raises(P.PannError):
        _static_gt(P, L, torch, idx, qd, 10, 128, 3001)      # k > n
    # the largest batch that fits still answers correctly
    nq_ok = (16 << 20) // (8 * k + 128 + 256) - 8
    qs = O.gen_points("uint8", nq_ok, 128, seed=3, stream=1)
    gi, gd = _static_gt(P, L, torch, idx, torch.from_numpy(qs).cuda(), nq_ok, 128, k)
    ref_i, ref_d = idx.groundtruth(qs, k=k)
    assert np.array_equal(gd, ref_d) and np.array_equal(gi, ref_i)
