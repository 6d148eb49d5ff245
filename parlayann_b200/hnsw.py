"""HNSW on the GPU behind the reference's surface (algorithms/HNSW/HNSW.hpp; python/builder.cpp:201-236 build_hnsw_index,
python/graph_index.cpp:103-109 load): build, save / load of the model file v3, batched search."""
import ctypes as C

import numpy as np

from . import _lib, fileio
from ._lib import check
from .hnsw_file import HnswModel, read_hnsw, write_hnsw
from .index import GraphIndex, _vp, _rstride


class HNSW:
    """ANN::HNSW<Desc_HNSW<T, Point>> over the rows of `points` (node i = point i)."""

    def __init__(self, points=None, data_type="float", metric="Euclidian", m=32, efc=128, m_l=0.36, alpha=1.0, device=0, synthetic=None):
        self.m, self.efc, self.m_l, self.alpha = int(m), int(efc), float(np.float32(m_l)), float(np.float32(alpha))
        self.index = GraphIndex(points, None, data_type, metric, max_degree=2 * self.m, device=device, synthetic=synthetic)
        self.n, self.d, self.data_type, self.metric = self.index.n, self.index.d, data_type, metric
        self._h = C.c_void_p()
        check(_lib.lib().pann_hnsw_create(C.byref(self._h), self.index._h, C.c_uint32(self.m), C.c_uint32(self.efc), C.c_float(self.m_l), C.c_float(self.alpha)))
        self._code = (0, 0)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            _lib.lib().pann_hnsw_destroy(self._h)
            self._h = C.c_void_p()
        if getattr(self, "index", None) is not None:
            self.index.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def build(self, levels=None):
        """HNSW(begin, end, dim, m_l, m, efc, alpha) (HNSW.hpp:746-824); returns (total, search, prune, reverse) seconds"""
        lv = None if levels is None else np.ascontiguousarray(levels, dtype=np.uint32)
        sec = (C.c_double * 4)()
        check(_lib.lib().pann_hnsw_build(self._h, _vp(lv), sec))
        self.build_seconds = tuple(sec)
        return self.build_seconds

    def model(self):
        sizes = (C.c_uint64 * 3)()
        check(_lib.lib().pann_hnsw_get_model(self._h, None, None, None, None, sizes))
        level = np.zeros(self.n, np.uint32); counts = np.zeros(sizes[0], np.uint64); edges = np.zeros(sizes[1], np.uint32)
        ent = np.zeros(sizes[2], np.uint32)
        check(_lib.lib().pann_hnsw_get_model(self._h, _vp(level), _vp(counts), _vp(edges), _vp(ent), sizes))
        return HnswModel(self.d, self.m_l, self.m, self.efc, self.alpha, level, counts, edges, ent, *self._code)

    def set_model(self, M):
        if M.n != self.n or M.m != self.m:
            raise ValueError("model and index disagree on n or m")
        if not np.array_equal(M.ids, np.arange(M.n, dtype=np.uint32)):
            raise ValueError("model files whose node ids are not the point ids 0..n-1 are not supported")
        self._code = (M.code_U, M.size_node)
        check(_lib.lib().pann_hnsw_set_model(self._h, _vp(M.level), _vp(M.counts), _vp(M.edges), _vp(M.entrance), C.c_uint64(len(M.entrance))))

    def save(self, path):
        write_hnsw(path, self.model())

    def search(self, queries, k=10, ef=50, limit=-1, ocap=None):
        """HNSW::search per row of `queries` -> (ids [nq, ocap], dists, cnt, cmps); cnt[i] = entries the reference returns
        (k plus ties with the k-th distance, HNSW.hpp:1594-1600)"""
        queries = np.ascontiguousarray(queries, dtype=fileio.np_dtype(self.data_type))
        nq = queries.shape[0]
        ocap = int(k if ocap is None else ocap)
        ids = np.empty((nq, ocap), np.uint32); dists = np.empty((nq, ocap), np.float32)
        cnt = np.empty(nq, np.uint32); cmps = np.empty(nq, np.uint32)
        check(_lib.lib().pann_hnsw_search(self._h, _vp(queries), C.c_uint64(nq), C.c_uint64(_rstride(queries)), C.c_uint32(k),
                                          C.c_uint32(ef), C.c_int64(limit), C.c_uint32(ocap), _vp(ids), _vp(dists), _vp(cnt), _vp(cmps)))
        return ids, dists, cnt, cmps

    def info(self):
        top, rows, nent = C.c_uint32(), C.c_uint64(), C.c_uint32()
        sec = (C.c_double * 4)()
        check(_lib.lib().pann_hnsw_info(self._h, C.byref(top), C.byref(rows), C.byref(nent), sec))
        return dict(top_level=top.value, upper_rows=rows.value, entrances=nent.value, seconds=tuple(sec))


def hnsw_levels(n, m_l):
    out = np.empty(n, np.uint32)
    check(_lib.lib().pann_hnsw_levels(C.c_uint64(n), C.c_float(m_l), _vp(out)))
    return out


def build_hnsw_index(metric, dtype, data_dir, index_dir, R, L, alpha, m_l=0.36, device=0):
    """python/wrapper.py:73-93 -> builder.cpp:201-236: HNSW over the points file, model saved to index_dir"""
    pts = fileio.read_bin(data_dir, dtype)
    h = HNSW(pts, dtype, metric, m=R, efc=L, m_l=m_l, alpha=alpha, device=device)
    h.build()
    h.save(index_dir)
    return h


def load_hnsw_index(metric, dtype, data_dir, index_dir, device=0):
    """GraphIndex(data_path, index_path, is_hnsw=true) (python/graph_index.cpp:103-109)"""
    pts = fileio.read_bin(data_dir, dtype)
    M = read_hnsw(index_dir)
    h = HNSW(pts, dtype, metric, m=M.m, efc=M.efc, m_l=M.m_l, alpha=M.alpha, device=device)
    h.set_model(M)
    return h
