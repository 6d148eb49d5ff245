"""ctypes binding of libpann_cuda.so (include/pann.h).  There is no CPU fallback: if the library
is missing, or a compute call is made without a CUDA device, this raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PANN_LIB_PATH") or os.path.join(_HERE, "libpann_cuda.so")  # the override is for A/B builds of the same ABI

DTYPE = {"float": 0, "uint8": 1, "int8": 2}
METRIC = {"Euclidian": 0, "mips": 1}
GEN_KIND = {"uint8": 0, "int8": 1, "float": 2, "float_norm": 3}


class QueryParams(C.Structure):
    """= parlayANN::QueryParams (reference: algorithms/utils/types.h:218-231)."""
    _fields_ = [("k", C.c_int64), ("beam", C.c_int64), ("cut", C.c_double), ("limit", C.c_int64),
                ("degree_limit", C.c_int64)]


class SweepResult(C.Structure):
    """= pann_sweep_result (include/pann.h): one row of the reference's search sweep (check_nn_recall.h:111-123)"""
    _fields_ = [("recall", C.c_double), ("avg_cmps", C.c_uint32), ("tail_cmps", C.c_uint32), ("avg_visited", C.c_uint32),
                ("tail_visited", C.c_uint32), ("device_ms", C.c_float), ("status", C.c_int)]


class PannError(RuntimeError):
    pass


# every symbol include/pann.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "pann_last_error", "pann_device_count", "pann_version", "pann_index_create", "pann_index_create_synthetic", "pann_index_create_synthetic_shard",
    "pann_index_destroy", "pann_index_set_graph", "pann_index_get_graph", "pann_index_get_points", "pann_index_info",
    "pann_index_last_timing", "pann_index_build_stats", "pann_index_last_bytes", "pann_search_batch", "pann_queries_upload", "pann_queries_synthetic",
    "pann_queries_download", "pann_search_resident", "pann_search_fetch", "pann_search_device_result", "pann_search_full", "pann_search_batch_starts", "pann_search_full_starts", "pann_search_rerank_batch", "pann_vamana_build", "pann_vamana_batch_insert",
    "pann_random_permutation", "pann_random_starts", "pann_batch_schedule", "pann_groundtruth", "pann_groundtruth_resident",
    "pann_groundtruth_resident_range", "pann_groundtruth_fetch", "pann_index_groundtruth_stats", "pann_groundtruth_static", "pann_groundtruth_static_limit", "pann_groundtruth_merge", "pann_groundtruth_device_result", "pann_groundtruth_merge_device",
    "pann_groundtruth_upload", "pann_search_sweep",
    "pann_hnsw_create", "pann_hnsw_destroy", "pann_hnsw_levels", "pann_hnsw_build", "pann_hnsw_set_model", "pann_hnsw_get_model",
    "pann_hnsw_search", "pann_hnsw_info", "pann_host_alloc", "pann_host_free", "pann_rows_normalize", "pann_rows_quant_params_u8", "pann_rows_quantize_u8",
    "pann_rows_quant_params_mips8", "pann_rows_quantize_mips8",
]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PannError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                            "(make -C parlayann_b200/csrc).  parlayann_b200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        L.pann_last_error.restype = C.c_char_p
        L.pann_version.restype = C.c_char_p
        L.pann_batch_schedule.restype = C.c_int64
        L.pann_index_destroy.restype = None
        L.pann_hnsw_destroy.restype = None
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise PannError(lib().pann_last_error().decode("utf-8", "replace"))


def device_count():
    n = C.c_int(0)
    rc = lib().pann_device_count(C.byref(n))
    return 0 if rc else n.value
