"""Drop-in for the reference's pybind module `_ParlayANNpy` (python/module.cpp:127-176): the same function and class names,
argument names and order, so that scripts written against the reference (python/scripts/sift_test.py: `import _ParlayANNpy as
pann; import wrapper as wp`) run unchanged with this directory on PYTHONPATH.  All compute happens in libpann_cuda.so on the GPU.

    PYTHONPATH=/path/to/repo:/path/to/repo/parlayann_b200/compat python sift_test.py

Covered: the Vamana builders and the six index classes (module.cpp:50-76), the HNSW builders (:123-126).  The HCNNG and
pyNNDescent builders (module.cpp:78-112) are outside this repository's scope and raise."""
import os
import sys
import types

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

import parlayann_b200 as _P  # noqa: E402
from parlayann_b200 import fileio as _fileio  # noqa: E402
from parlayann_b200 import hnsw as _hnsw  # noqa: E402
from parlayann_b200 import index as _index  # noqa: E402

__version__ = "parlayann_b200"

defaults = types.ModuleType("_ParlayANNpy.defaults")   # module.cpp:139-148
defaults.METRIC = "Euclidian"
defaults.ALPHA = 1.2
defaults.GRAPH_DEGREE = 64
defaults.BEAMWIDTH = 128
sys.modules["_ParlayANNpy.defaults"] = defaults

_DTYPES = {"float": "Float", "uint8": "UInt8", "int8": "Int8"}
_METRICS = {"Euclidian": "Euclidian", "mips": "Mips"}


class _GraphIndex:
    """GraphIndex<T, Point> (python/graph_index.cpp:48-337).  NOTE the constructor's keyword names: the reference binds
    init<std::string&, std::string&, bool> as ("index_path", "data_path", "hnsw") onto the C++ constructor
    (data_path, index_path, is_hnsw) — the FIRST argument is the points file, the second the graph / model file, and
    wrapper.load_index passes (data_dir, index_dir) positionally (module.cpp:66-67, graph_index.cpp:82, wrapper.py:96-116)."""
    _dtype = None
    _metric = None

    def __init__(self, index_path, data_path, hnsw=False):
        points_file, graph_file = index_path, data_path
        self._is_hnsw = bool(hnsw)
        if self._is_hnsw:
            self._h = _hnsw.load_hnsw_index(self._metric, self._dtype, points_file, graph_file)
            self._g = self._h.index
        else:
            self._g = _index.load_index(self._metric, self._dtype, points_file, graph_file)

    def batch_search(self, queries, knn, beam_width, quant=False, visit_limit=-1):
        if self._is_hnsw:   # the dispatch the reference has commented out (graph_index.cpp:122-136): HNSW::search(q, k, ef, limit)
            ids, dists, _, _ = self._h.search(np.asarray(queries), k=knn, ef=beam_width, limit=visit_limit)
            return ids, dists
        return self._g.batch_search(queries, knn, beam_width, quant, visit_limit)

    def single_search(self, q, knn, beam_width, quant=False, visit_limit=-1):
        return self.batch_search(np.asarray(q)[None, :], knn, beam_width, quant, visit_limit)[0][0]

    def batch_search_from_string(self, queries, knn, beam_width, quant=False, visit_limit=-1):
        return self.batch_search(_fileio.read_bin(queries, self._dtype), knn, beam_width, quant, visit_limit)

    def check_recall(self, queries_file, graph_file, neighbors, k):
        """graph_file is the ground-truth .ibin (graph_index.cpp:264)"""
        return self._g.check_recall(queries_file, graph_file, neighbors, k)


def _make(dtype, metric):
    tag = _DTYPES[dtype] + _METRICS[metric]
    low = f"{dtype}_{_METRICS[metric].lower()}"

    cls = type(tag + "Index", (_GraphIndex,), {"_dtype": dtype, "_metric": metric})

    def build_vamana(distance_metric, data_file_path, index_output_path, graph_degree, beam_width, alpha, two_pass):
        _index.build_vamana_index(distance_metric, dtype, data_file_path, index_output_path, graph_degree, beam_width, alpha, two_pass).close()

    def build_hnsw(distance_metric, data_file_path, index_output_path, graph_degree, efc, m_l, alpha):
        _hnsw.build_hnsw_index(distance_metric, dtype, data_file_path, index_output_path, graph_degree, efc, alpha, m_l=m_l).close()

    def unsupported(*a, **k):
        raise NotImplementedError("the HCNNG / pyNNDescent builders are outside this repository's scope (SURVEY.md §8f.4)")

    g = globals()
    g[tag + "Index"] = cls
    g[f"build_vamana_{low}_index"] = build_vamana
    g[f"build_hnsw_{low}_index"] = build_hnsw
    g[f"build_hcnng_{low}_index"] = unsupported
    g[f"build_pynndescent_{low}_index"] = unsupported


for _d in _DTYPES:
    for _m in _METRICS:
        _make(_d, _m)
