"""Stand-in for the reference's python/wrapper.py (same function names and argument order), dispatching on (metric, dtype) to
the `_ParlayANNpy` names next to it."""
import _ParlayANNpy as _m
from _ParlayANNpy import *  # noqa: F401,F403

_T = {"float": "Float", "uint8": "UInt8", "int8": "Int8"}
_M = {"Euclidian": "Euclidian", "mips": "Mips"}


def _pick(kind, metric, dtype):
    if metric not in _M:
        raise Exception("Invalid metric " + str(metric))
    if dtype not in _T:
        raise Exception("Invalid data type " + str(dtype))
    return getattr(_m, f"build_{kind}_{dtype}_{_M[metric].lower()}_index")


def build_vamana_index(metric, dtype, data_dir, index_dir, R, L, alpha, two_pass):
    return _pick("vamana", metric, dtype)(metric, data_dir, index_dir, R, L, alpha, two_pass)


def build_hcnng_index(metric, dtype, data_dir, index_dir, mst_deg, num_clusters, cluster_size):
    return _pick("hcnng", metric, dtype)(metric, data_dir, index_dir, mst_deg, num_clusters, cluster_size)


def build_pynndescent_index(metric, dtype, data_dir, index_dir, max_deg, num_clusters, cluster_size, alpha, delta):
    return _pick("pynndescent", metric, dtype)(metric, data_dir, index_dir, max_deg, num_clusters, cluster_size, alpha, delta)


def build_hnsw_index(metric, dtype, data_dir, index_dir, R, L, alpha, two_pass):
    """the reference forwards (R, L, alpha, two_pass) into (graph_degree, efc, m_l, alpha) (wrapper.py:73-93, module.cpp:123-126)"""
    return _pick("hnsw", metric, dtype)(metric, data_dir, index_dir, R, L, alpha, two_pass)


def load_index(metric, dtype, data_dir, index_dir, hnsw=False):
    if metric not in _M:
        raise Exception("Invalid metric")
    if dtype not in _T:
        raise Exception("Invalid data type")
    return getattr(_m, _T[dtype] + _M[metric] + "Index")(data_dir, index_dir, hnsw)
