"""Host-side mirror of the reference's Python surface for the hot path, on top of the C ABI.

Reference names kept (python/wrapper.py:3-116, python/module.cpp:59-76, python/graph_index.cpp:48-337):
``build_vamana_index``, ``load_index``, and index objects with ``batch_search``, ``single_search``,
``batch_search_from_string``, ``check_recall``.  All compute happens in libpann_cuda.so on the GPU.
"""
import ctypes as C
import numpy as np

from . import _lib, fileio
from ._lib import QueryParams, PannError, check


def _vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _rstride(a):
    """bytes between the rows of a C-contiguous 2-D array (numpy reports an arbitrary stride, even 0, for an axis of length 1)"""
    return a.shape[1] * a.itemsize if a.ndim == 2 else a.strides[0]


def quant_params_u8(points):
    """Euclidian_Point<uint8_t>::generate_parameters (euclidian_point.h:211-237) + parameters(min, max, dims) (:104-108):
    per-row minima / maxima start at 0, all-integer non-negative data below 256 maps unscaled.  Returns (slope, offset)."""
    p = np.asarray(points, dtype=np.float32)
    min_val = np.float32(min(0.0, float(p.min()))) if p.size else np.float32(0)
    max_val = np.float32(max(0.0, float(p.max()))) if p.size else np.float32(0)
    if p.size == 0 or bool(((p >= 0) & (p == np.trunc(p))).all()):
        if max_val < 256:
            max_val = np.float32(255)
        min_val = np.float32(0)
    slope = np.float32(255) / np.float32(max_val - min_val)
    v = float(np.float32(min_val * slope))
    offset = int(np.trunc(v) + (np.sign(v) if abs(v - np.trunc(v)) >= 0.5 else 0))   # round(), halves away from zero
    return slope, offset


def quantize_u8(points, slope, offset):
    """Euclidian_Point<uint8_t>::translate_point (euclidian_point.h:183-209): (int64)std::round(x * slope) - offset,
    clamped to [0, 255]; the product is a float product, std::round rounds halves away from zero."""
    p = np.asarray(points, dtype=np.float32)
    if np.float32(slope) == np.float32(1.0) and offset == 0:
        return p.astype(np.uint8)
    v = p * np.float32(slope)                       # float32 product
    t = np.trunc(v)
    r = t + np.where(np.abs(v - t) >= np.float32(0.5), np.sign(v), np.float32(0))   # v - t is exact
    return np.clip(r.astype(np.int64) - int(offset), 0, 255).astype(np.uint8)


def normalize_rows(points):
    """Mips_Point<float>::normalize over every row (mips_point.h:115-124): the float squares are summed into a double in
    element order, 1/sqrt is rounded to float, the row is scaled by float products.  Returns a new float32 array."""
    p = np.array(points, dtype=np.float32, order="C")
    norm = np.zeros(p.shape[0], np.float64)
    for j in range(p.shape[1]):                     # element order is part of the result
        norm += (p[:, j] * p[:, j]).astype(np.float64)
    norm = np.sqrt(norm)
    norm[norm == 0] = 1.0
    p *= (1.0 / norm).astype(np.float32)[:, None]
    return p


def quant_params_mips8(points, trim=True):
    """Quantized_Mips_Point<8,trim>::generate_parameters (mips_point.h:437-462): bound = max(max_val, -min_val) over all
    values, after dropping 1e-4 of each tail when trim is set (graph_index.cpp and builder.cpp use trim).  The lower index
    is a float product, the upper one a double product, as in the reference.  Returns max_val (float32)."""
    v = np.asarray(points, dtype=np.float32).reshape(-1)
    ln = v.size
    if ln == 0:
        return np.float32(1.0)
    if trim:
        cutoff = np.float32(.0001)
        lo = int(np.float32(cutoff * np.float32(ln)))
        hi = int((1.0 - float(cutoff)) * (ln - 1))
        part = np.partition(v, sorted({lo, hi}))
        min_val, max_val = part[lo], part[hi]
    else:
        min_val, max_val = v.min(), v.max()
    return np.float32(max(max_val, -min_val))


def quantize_mips8(points, max_val):
    """Quantized_Mips_Point<8>::translate_point (mips_point.h:421-434): scale = 127 / max_val (float), values beyond
    +-max_val cap at +-127, the rest std::round(x * scale) with halves away from zero.  Returns int8."""
    p = np.asarray(points, dtype=np.float32)
    mv = np.float32(max_val)
    scale = np.float32(127) / mv
    v = p * scale
    t = np.trunc(v)
    r = t + np.where(np.abs(v - t) >= np.float32(0.5), np.sign(v), np.float32(0))
    r = np.where(p < -mv, np.float32(-127), np.where(p > mv, np.float32(127), r))
    return r.astype(np.int32).astype(np.int8)

# ---- the same conversions on the device (csrc/quantize.cu): what the product path uses; the numpy versions above are the
# host-side mirror the CPU tests pin against the reference's bytes (tests/test_quantized_cpu.py)
def _f32_rows(points):
    return np.ascontiguousarray(points, dtype=np.float32)    # _rstride assumes dense rows; the C call itself takes any row stride


def normalize_rows_device(points, device=0):
    p = np.array(points, dtype=np.float32, order="C")
    check(_lib.lib().pann_rows_normalize(device, _vp(p), C.c_uint64(p.shape[0]), C.c_uint32(p.shape[1]), C.c_uint64(_rstride(p))))
    return p


def quant_params_u8_device(points, device=0):
    p = _f32_rows(points)
    slope, offset = C.c_float(0), C.c_int64(0)
    check(_lib.lib().pann_rows_quant_params_u8(device, _vp(p), C.c_uint64(p.shape[0]), C.c_uint32(p.shape[1]), C.c_uint64(_rstride(p)),
                                               C.byref(slope), C.byref(offset)))
    return np.float32(slope.value), int(offset.value)


def quantize_u8_device(points, slope, offset, device=0):
    p = _f32_rows(points)
    out = np.empty(p.shape, np.uint8)
    check(_lib.lib().pann_rows_quantize_u8(device, _vp(p), C.c_uint64(p.shape[0]), C.c_uint32(p.shape[1]), C.c_uint64(_rstride(p)),
                                           C.c_float(float(np.float32(slope))), C.c_int64(int(offset)), _vp(out), C.c_uint64(p.shape[1])))
    return out


def quant_params_mips8_device(points, trim=True, device=0):
    p = _f32_rows(points)
    mv = C.c_float(0)
    check(_lib.lib().pann_rows_quant_params_mips8(device, _vp(p), C.c_uint64(p.shape[0]), C.c_uint32(p.shape[1]), C.c_uint64(_rstride(p)),
                                                  C.c_int(1 if trim else 0), C.byref(mv)))
    return np.float32(mv.value)


def quantize_mips8_device(points, max_val, device=0):
    p = _f32_rows(points)
    out = np.empty(p.shape, np.int8)
    check(_lib.lib().pann_rows_quantize_mips8(device, _vp(p), C.c_uint64(p.shape[0]), C.c_uint32(p.shape[1]), C.c_uint64(_rstride(p)),
                                              C.c_float(float(np.float32(max_val))), _vp(out), C.c_uint64(p.shape[1])))
    return out


class _PinnedBlock:
    """page-locked host memory from the library (cudaHostAlloc), freed with the last numpy view of it"""

    def __init__(self, nbytes):
        self.p = C.c_void_p()
        check(_lib.lib().pann_host_alloc(C.byref(self.p), C.c_size_t(max(1, nbytes))))
        self.nbytes = nbytes

    def __del__(self):
        try:
            if self.p:
                _lib.lib().pann_host_free(self.p)
        except Exception:
            pass


def pinned_empty(shape, dtype):
    """np.empty in page-locked memory: copies to and from the device run at full PCIe speed and can overlap kernels"""
    dt = np.dtype(dtype)
    n = int(np.prod(shape)) * dt.itemsize
    blk = _PinnedBlock(n)
    buf = (C.c_uint8 * max(1, n)).from_address(blk.p.value)
    buf._pann_block = blk      # the numpy array keeps `buf` alive (its base), `buf` keeps the block: freed with the last view
    return np.frombuffer(buf, dtype=dt, count=int(np.prod(shape))).reshape(shape)


class GraphIndex:
    """PointRange + Graph resident on one B200 (reference: python/graph_index.cpp:48-118)."""

    def __init__(self, points=None, graph=None, data_type="uint8", metric="Euclidian", max_degree=None, device=0,
                 synthetic=None):
        self.data_type, self.metric, self.device = data_type, metric, device
        self._h = C.c_void_p()
        L = _lib.lib()
        if synthetic is not None:  # dict(n, d, kind, seed, stream, r, nc): points generated on the device
            s = synthetic
            self.n, self.d = int(s["n"]), int(s["d"])
            self.max_degree = int(max_degree)
            check(L.pann_index_create_synthetic_shard(C.byref(self._h), device, _lib.DTYPE[data_type], _lib.METRIC[metric],
                                                C.c_uint64(s.get("row0", 0)), C.c_uint64(self.n), C.c_uint32(self.d), C.c_uint32(self.max_degree),
                                                _lib.GEN_KIND[s.get("kind", data_type)], C.c_uint64(s.get("seed", 1234)),
                                                C.c_uint64(s.get("stream", 0)), C.c_uint32(s.get("r", 24)),
                                                C.c_uint32(s.get("nc", 256))))
        else:
            points = np.ascontiguousarray(points, dtype=fileio.np_dtype(data_type))
            self.n, self.d = points.shape
            if graph is not None:
                graph = np.ascontiguousarray(graph, dtype=np.uint32)
                if graph.shape[0] != self.n:
                    raise ValueError("graph and points disagree on n")
                max_degree = graph.shape[1] - 1
            self.max_degree = int(max_degree)
            check(L.pann_index_create(C.byref(self._h), device, _lib.DTYPE[data_type], _lib.METRIC[metric],
                                      C.c_uint64(self.n), C.c_uint32(self.d), C.c_uint64(_rstride(points)), _vp(points),
                                      C.c_uint32(self.max_degree), _vp(graph)))
        self.start_point = 0  # knn_index::set_start (vamana/index.h:148)

    def close(self):
        if getattr(self, "_qidx", None) is not None:
            self._qidx.close()
            self._qidx = None
        if getattr(self, "_h", None) is not None and self._h:
            _lib.lib().pann_index_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- graph / points access
    def set_graph(self, graph):
        graph = np.ascontiguousarray(graph, dtype=np.uint32)
        assert graph.shape == (self.n, self.max_degree + 1)
        check(_lib.lib().pann_index_set_graph(self._h, _vp(graph)))

    def graph(self):
        g = np.empty((self.n, self.max_degree + 1), dtype=np.uint32)
        check(_lib.lib().pann_index_get_graph(self._h, _vp(g)))
        return g

    def points(self, row0=0, nrows=None):
        nrows = self.n - row0 if nrows is None else nrows
        out = np.empty((nrows, self.d), dtype=fileio.np_dtype(self.data_type))
        check(_lib.lib().pann_index_get_points(self._h, C.c_uint64(row0), C.c_uint64(nrows), _vp(out), C.c_uint64(_rstride(out))))
        return out

    def last_timing(self):
        ms, nl = C.c_float(0), C.c_int(0)
        check(_lib.lib().pann_index_last_timing(self._h, C.byref(ms), C.byref(nl)))
        return ms.value, nl.value

    def build_stats(self):
        """accounting of the last build (pann_index_build_stats): dict of algorithmic bytes and call counts"""
        v = [C.c_uint64(0) for _ in range(5)]
        check(_lib.lib().pann_index_build_stats(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("algorithmic_bytes", "search_bytes", "prune_bytes", "prune_calls", "pair_distances"), (x.value for x in v)))

    def last_bytes(self):
        b = C.c_uint64(0)
        check(_lib.lib().pann_index_last_bytes(self._h, C.byref(b)))
        return b.value

    # ---- search (qsearchAll, beamSearch.h:537-565)
    def _qp(self, k, beam, cut, limit, degree_limit):
        return QueryParams(int(k), int(beam), float(cut), -1 if limit is None else int(limit),
                           -1 if degree_limit is None else int(degree_limit))

    def search(self, queries, k=10, beam=10, cut=1.35, limit=None, degree_limit=None, starts=None, out=None, nstarts_q=None):
        """One beam search per row of `queries` (host array).  Returns ids, dists, visited, cmps."""
        queries = np.ascontiguousarray(queries, dtype=fileio.np_dtype(self.data_type))
        nq = queries.shape[0]
        st = np.ascontiguousarray([self.start_point] if starts is None else starts, dtype=np.uint32)
        if out is None:
            out = (np.empty((nq, k), np.uint32), np.empty((nq, k), np.float32), np.empty(nq, np.uint32), np.empty(nq, np.uint32))
        qp = self._qp(k, beam, cut, limit, degree_limit)
        if st.ndim == 2:   # per-query start lists [nq, stride] (beamSearchRandom, beamSearch.h:309-351; HNSW entry points)
            if st.shape[0] != nq:
                raise ValueError("per-query starts need one row per query")
            nsq = None if nstarts_q is None else np.ascontiguousarray(nstarts_q, dtype=np.uint32)
            check(_lib.lib().pann_search_batch_starts(self._h, _vp(queries), C.c_uint64(nq), C.c_uint64(_rstride(queries)), _vp(st),
                                                      C.c_uint32(st.shape[1]), _vp(nsq), C.byref(qp), _vp(out[0]), _vp(out[1]), _vp(out[2]), _vp(out[3])))
            return out
        check(_lib.lib().pann_search_batch(self._h, _vp(queries), C.c_uint64(nq), C.c_uint64(_rstride(queries)), _vp(st),
                                           C.c_uint32(len(st)), C.byref(qp), _vp(out[0]), _vp(out[1]), _vp(out[2]), _vp(out[3])))
        return out

    def upload_queries(self, queries):
        queries = np.ascontiguousarray(queries, dtype=fileio.np_dtype(self.data_type))
        check(_lib.lib().pann_queries_upload(self._h, _vp(queries), C.c_uint64(queries.shape[0]), C.c_uint64(_rstride(queries))))
        self.nq = queries.shape[0]

    def synthetic_queries(self, nq, kind=None, seed=1234, stream=1, r=24, nc=256):
        check(_lib.lib().pann_queries_synthetic(self._h, C.c_uint64(nq), _lib.GEN_KIND[kind or self.data_type], C.c_uint64(seed),
                                                C.c_uint64(stream), C.c_uint32(r), C.c_uint32(nc)))
        self.nq = nq

    def download_queries(self):
        out = np.empty((self.nq, self.d), dtype=fileio.np_dtype(self.data_type))
        check(_lib.lib().pann_queries_download(self._h, _vp(out), C.c_uint64(_rstride(out))))
        return out

    def search_resident(self, k=10, beam=10, cut=1.35, limit=None, degree_limit=None, starts=None):
        """Search the resident query set; results stay on the device (fetch_results)."""
        st = np.asarray([self.start_point] if starts is None else starts, dtype=np.uint32)
        qp = self._qp(k, beam, cut, limit, degree_limit)
        check(_lib.lib().pann_search_resident(self._h, _vp(st), C.c_uint32(len(st)), C.byref(qp)))
        self._res_k = k

    def fetch_results(self):
        nq, k = self.nq, self._res_k
        out = (np.empty((nq, k), np.uint32), np.empty((nq, k), np.float32), np.empty(nq, np.uint32), np.empty(nq, np.uint32))
        check(_lib.lib().pann_search_fetch(self._h, _vp(out[0]), _vp(out[1]), _vp(out[2]), _vp(out[3])))
        return out

    def search_full(self, queries=None, self_ids=None, k=0, beam=10, cut=1.35, limit=None, degree_limit=None, starts=None,
                    vcap=1024, nstarts_q=None):
        """beam_search (beamSearch.h:217-223): whole frontier + whole visited list per query."""
        if self_ids is not None:
            self_ids = np.ascontiguousarray(self_ids, dtype=np.uint32)
            nq = len(self_ids)
            qptr, qstride = None, 0
        else:
            queries = np.ascontiguousarray(queries, dtype=fileio.np_dtype(self.data_type))
            nq = queries.shape[0]
            qptr, qstride = _vp(queries), _rstride(queries)
        st = np.ascontiguousarray([self.start_point] if starts is None else starts, dtype=np.uint32)
        qp = self._qp(k, beam, cut, limit, degree_limit)
        fcap = int(max(beam, st.shape[-1]))
        r = dict(fr_ids=np.zeros((nq, fcap), np.uint32), fr_dists=np.zeros((nq, fcap), np.float32), nfr=np.zeros(nq, np.uint32),
                 vis_ids=np.zeros((nq, vcap), np.uint32), vis_dists=np.zeros((nq, vcap), np.float32), nvis=np.zeros(nq, np.uint32),
                 cmps=np.zeros(nq, np.uint32))
        if st.ndim == 2:
            nsq = None if nstarts_q is None else np.ascontiguousarray(nstarts_q, dtype=np.uint32)
            check(_lib.lib().pann_search_full_starts(self._h, qptr, C.c_uint64(nq), C.c_uint64(qstride), _vp(self_ids), _vp(st),
                                                     C.c_uint32(st.shape[1]), _vp(nsq), C.byref(qp), C.c_uint32(fcap), _vp(r["fr_ids"]),
                                                     _vp(r["fr_dists"]), _vp(r["nfr"]), C.c_uint32(vcap), _vp(r["vis_ids"]),
                                                     _vp(r["vis_dists"]), _vp(r["nvis"]), _vp(r["cmps"])))
            return r
        check(_lib.lib().pann_search_full(self._h, qptr, C.c_uint64(nq), C.c_uint64(qstride), _vp(self_ids), _vp(st),
                                          C.c_uint32(len(st)), C.byref(qp), C.c_uint32(fcap), _vp(r["fr_ids"]), _vp(r["fr_dists"]),
                                          _vp(r["nfr"]), C.c_uint32(vcap), _vp(r["vis_ids"]), _vp(r["vis_dists"]), _vp(r["nvis"]),
                                          _vp(r["cmps"])))
        return r

    # ---- build (knn_index::build_index, vamana/index.h:150-186)
    def build(self, R=None, L=128, alpha=1.2, num_passes=1, sort_neighbors=True, stats=False):
        R = self.max_degree if R is None else R
        if R != self.max_degree:
            raise ValueError("R must equal the index's max degree (BuildParams::max_degree, types.h:210-214)")
        sv = np.zeros(self.n, np.uint32) if stats else None
        sd = np.zeros(self.n, np.uint32) if stats else None
        ph = (C.c_double * 3)()
        check(_lib.lib().pann_vamana_build(self._h, C.c_uint32(R), C.c_uint32(L), C.c_double(alpha), int(num_passes),
                                           int(bool(sort_neighbors)), _vp(sv), _vp(sd), ph))
        self.build_phase_seconds = tuple(ph)
        return (sv, sd) if stats else None

    def batch_insert(self, inserts, L=128, alpha=1.2, random_order=False, max_fraction=0.02, stats=False):
        """knn_index::batch_insert (vamana/index.h:188-316): insert the listed vertices into the graph the index holds"""
        ins = np.ascontiguousarray(inserts, dtype=np.uint32)
        sv = np.zeros(self.n, np.uint32) if stats else None
        sd = np.zeros(self.n, np.uint32) if stats else None
        ph = (C.c_double * 3)()
        check(_lib.lib().pann_vamana_batch_insert(self._h, _vp(ins), C.c_uint64(len(ins)), C.c_uint32(self.max_degree), C.c_uint32(L),
                                                  C.c_double(alpha), int(bool(random_order)), C.c_double(max_fraction), _vp(sv), _vp(sd), ph))
        self.build_phase_seconds = tuple(ph)
        return (sv, sd) if stats else None

    # ---- ground truth (compute_groundtruth.cpp:22-100)
    def groundtruth(self, queries, k=100, id_offset=0):
        queries = np.ascontiguousarray(queries, dtype=fileio.np_dtype(self.data_type))
        nq = queries.shape[0]
        ids = np.empty((nq, k), np.int32)
        dists = np.empty((nq, k), np.float32)
        check(_lib.lib().pann_groundtruth(self._h, _vp(queries), C.c_uint64(nq), C.c_uint64(_rstride(queries)), C.c_uint32(k),
                                          C.c_int64(id_offset), _vp(ids), _vp(dists)))
        return ids, dists

    def groundtruth_resident(self, k=100, id_offset=0):
        check(_lib.lib().pann_groundtruth_resident(self._h, C.c_uint32(k), C.c_int64(id_offset)))
        self._gt_k = k

    def groundtruth_range(self, k=100, id_offset=0, row_lo=0, row_hi=0, keep_state=False, d_tau_cap=0, seen_rows=0):
        """one leg of a sharded ground-truth job (pann_groundtruth_resident_range); d_tau_cap is a DEVICE pointer or 0"""
        check(_lib.lib().pann_groundtruth_resident_range(self._h, C.c_uint32(k), C.c_int64(id_offset), C.c_uint64(row_lo), C.c_uint64(row_hi),
                                                         C.c_int(1 if keep_state else 0), C.c_void_p(d_tau_cap or None), C.c_uint64(seen_rows)))
        self._gt_k = k

    def groundtruth_stats(self):
        """(launches of the tcgen05 kernel in the last ground-truth call, cumulative chunks retaken on the CUDA cores)"""
        a, b = C.c_int(0), C.c_int(0)
        check(_lib.lib().pann_index_groundtruth_stats(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def upload_groundtruth(self, gt_ids, gt_dists):
        """ground truth of the resident queries from a .ibin (fileio.read_ibin) for the device-side sweep"""
        gi = np.ascontiguousarray(gt_ids, dtype=np.int32); gd = np.ascontiguousarray(gt_dists, dtype=np.float32)
        check(_lib.lib().pann_groundtruth_upload(self._h, _vp(gi), _vp(gd), C.c_uint64(gi.shape[0]), C.c_uint32(gi.shape[1])))
        self._gt_k = gi.shape[1]

    def fetch_groundtruth(self):
        ids = np.empty((self.nq, self._gt_k), np.int32)
        dists = np.empty((self.nq, self._gt_k), np.float32)
        check(_lib.lib().pann_groundtruth_fetch(self._h, _vp(ids), _vp(dists)))
        return ids, dists

    # ---- quantised search + exact re-rank (beam_search_rerank with differing byte widths, beamSearch.h:390-454)
    def search_rerank(self, qindex, queries, qqueries, k=10, beam=10, cut=1.35, limit=None, degree_limit=None, starts=None,
                      rerank_factor=100):
        """`self` holds the exact points; `qindex` (same device) the 1-byte copy and the graph.  The search runs on qindex
        with `qqueries`; the frontier is re-ranked with exact distances to `queries`.  Returns ids, dists, visited, cmps."""
        queries = np.ascontiguousarray(queries, dtype=fileio.np_dtype(self.data_type))
        qqueries = np.ascontiguousarray(qqueries, dtype=fileio.np_dtype(qindex.data_type))
        nq = queries.shape[0]
        st = np.asarray([qindex.start_point] if starts is None else starts, dtype=np.uint32)
        out = (np.empty((nq, k), np.uint32), np.empty((nq, k), np.float32), np.empty(nq, np.uint32), np.empty(nq, np.uint32))
        qp = qindex._qp(k, beam, cut, limit, degree_limit)
        check(_lib.lib().pann_search_rerank_batch(qindex._h, self._h, _vp(qqueries), C.c_uint64(_rstride(qqueries)),
                                                  _vp(queries), C.c_uint64(_rstride(queries)), C.c_uint64(nq), _vp(st),
                                                  C.c_uint32(len(st)), C.byref(qp), C.c_int64(rerank_factor),
                                                  _vp(out[0]), _vp(out[1]), _vp(out[2]), _vp(out[3])))
        return out

    def _quantized(self):
        """the 1-byte copy graph_index.cpp keeps for 4-byte data (:57-95): Euclidian_Point<uint8_t> for metric points,
        Quantized_Mips_Point<8,true> for MIPS (whose exact rows were normalised at load time); built once, on first use"""
        if getattr(self, "_qidx", None) is None:
            if self.data_type != "float":
                raise NotImplementedError("quant=True is built for float data")
            if self.d > (800 if self.metric == "Euclidian" else 200):
                raise NotImplementedError("the second-level filter copies the reference adds above 800 (Euclidean) / 200 (MIPS) dimensions (graph_index.cpp:91-95, use_filtering) are not built")
            pts = self.points()
            if self.metric == "Euclidian":
                self._qparams = quant_params_u8_device(pts, self.device)
                self._qidx = GraphIndex(quantize_u8_device(pts, *self._qparams, device=self.device), self.graph(), "uint8", "Euclidian", device=self.device)
            else:
                self._qparams = (quant_params_mips8_device(pts, trim=True, device=self.device),)
                self._qidx = GraphIndex(quantize_mips8_device(pts, *self._qparams, device=self.device), self.graph(), "int8", "mips", device=self.device)
        return self._qidx

    # ---- the reference's Python methods (python/graph_index.cpp:192-305)
    def batch_search(self, queries, knn, beam_width, quant=False, visit_limit=-1):
        dl = min(self.max_degree, 3 * visit_limit)  # graph_index.cpp:198
        if quant and fileio.np_dtype(self.data_type)().itemsize >= 4:   # use_quantization (graph_index.cpp:60)
            qi = self._quantized()
            queries = np.ascontiguousarray(queries, dtype=np.float32)
            if self.metric != "Euclidian":            # graph_index.cpp:166-181: q.normalize(), Quantized_Mips_Point copy of it
                queries = normalize_rows_device(queries, self.device)
                qq = quantize_mips8_device(queries, *self._qparams, device=self.device)
            elif self._qparams[0] == np.float32(1.0):   # graph_index.cpp:146-150: plain search of the 1-byte copy
                ids, dists, _, _ = qi.search(queries.astype(np.uint8), k=knn, beam=beam_width, cut=1.35, limit=max(visit_limit, 0), degree_limit=max(dl, 0))
                return ids, dists
            else:
                qq = quantize_u8_device(queries, *self._qparams, device=self.device)
            ids, dists, _, _ = self.search_rerank(qi, queries, qq, k=knn, beam=beam_width, cut=1.35,
                                                  limit=max(visit_limit, 0), degree_limit=max(dl, 0))
            return ids, dists
        # the reference passes visit_limit straight into QueryParams: -1 means "visit nothing"
        ids, dists, _, _ = self.search(queries, k=knn, beam=beam_width, cut=1.35, limit=max(visit_limit, 0), degree_limit=max(dl, 0))
        return ids, dists

    def single_search(self, q, knn, beam_width, quant=False, visit_limit=-1):
        return self.batch_search(np.asarray(q)[None, :], knn, beam_width, quant, visit_limit)[0][0]

    def batch_search_from_string(self, queries_file, knn, beam_width, quant=False, visit_limit=-1):
        return self.batch_search(fileio.read_bin(queries_file, self.data_type), knn, beam_width, quant, visit_limit)

    def check_recall(self, queries_file, gt_file, neighbors, k):
        from .harness import recall_with_ties
        gt_ids, gt_dists = fileio.read_ibin(gt_file)
        r = recall_with_ties(neighbors, gt_ids, gt_dists, k)
        print(f"Recall: {r:.6g}")
        return r


def groundtruth_merge(ids_list, dists_list, device=0):
    """k-way merge of per-shard ground truth ([nq,k] each) under the (dist,id) order."""
    ids = np.ascontiguousarray(np.stack(ids_list), dtype=np.int32)
    dists = np.ascontiguousarray(np.stack(dists_list), dtype=np.float32)
    s, nq, k = ids.shape
    oi = np.empty((nq, k), np.int32)
    od = np.empty((nq, k), np.float32)
    check(_lib.lib().pann_groundtruth_merge(device, C.c_uint32(s), C.c_uint64(nq), C.c_uint32(k), _vp(ids), _vp(dists), _vp(oi), _vp(od)))
    return oi, od


def random_permutation(n):
    out = np.empty(n, np.int32)
    check(_lib.lib().pann_random_permutation(C.c_int64(n), _vp(out)))
    return out


def random_starts(nq, n):
    """beamSearchRandom's per-query start points (beamSearch.h:327-332)"""
    out = np.empty(nq, np.uint32)
    check(_lib.lib().pann_random_starts(C.c_uint64(nq), C.c_uint64(n), _vp(out)))
    return out


def batch_schedule(n, m=None):
    m = n if m is None else m
    lo = np.zeros(8192, np.int64)
    hi = np.zeros(8192, np.int64)
    nb = _lib.lib().pann_batch_schedule(C.c_uint64(n), C.c_uint64(m), _vp(lo), _vp(hi), C.c_int64(8192))
    return [(int(lo[i]), int(hi[i])) for i in range(nb)]


# ---- module-level names of the reference's wrapper (python/wrapper.py)
class defaults:  # python/module.cpp:142-148
    METRIC = "Euclidian"
    ALPHA = 1.2
    GRAPH_DEGREE = 64
    BEAMWIDTH = 128


def build_vamana_index(metric, dtype, data_dir, index_dir, R, L, alpha, two_pass, device=0):
    """python/wrapper.py:3-24 -> builder.cpp:36-95.  1-byte Euclidean types build directly on the points
    (builder.cpp:88-94); 4-byte types build on a 1-byte copy, as the reference's builder does, and the graph file holds the
    graph of that copy: Euclidian_Point<uint8_t> for Euclidean data (builder.cpp:60-72); float MIPS data is normalised
    first, alpha is overridden (1.0 below 100 dimensions, 0.98 up to 200; builder.cpp:45-54) and the copy is
    Quantized_Mips_Point<8,true> (builder.cpp:73-86), whose distance is the int8 inner product, so the build runs the int8
    MIPS kernels over those bytes (pinned against the reference in tests/test_quantized_cpu.py).  1-byte MIPS types would
    be normalised IN their integer type by the reference (rows collapse to 0/1 values): that raises here."""
    if metric not in _lib.METRIC:
        raise Exception("Invalid metric " + str(metric))
    if dtype not in _lib.DTYPE:
        raise Exception("Invalid data type " + str(dtype))
    wide = fileio.np_dtype(dtype)().itemsize > 1
    if metric != "Euclidian" and not wide:
        raise NotImplementedError("build_vamana_index for 1-byte MIPS data: the reference normalises the integer rows in place (builder.cpp:45-49), which is not reproduced")
    pts = fileio.read_bin(data_dir, dtype)
    alpha = float(np.float32(alpha))                   # builder.cpp takes alpha as a float
    if metric != "Euclidian":
        print("normalizing")
        pts = normalize_rows_device(pts, device)
        if pts.shape[1] <= 200:
            alpha = float(np.float32(1.0 if pts.shape[1] < 100 else .98))
        max_val = quant_params_mips8_device(pts, trim=True, device=device)
        print(f"mips scalar quantization to 8 bits: trimmed to max = {max_val}")
        idx = GraphIndex(quantize_mips8_device(pts, max_val, device=device), None, "int8", metric, max_degree=R, device=device)
    elif wide:
        slope, offset = quant_params_u8_device(pts, device)
        print(f"scalar quantization: slope = {slope}, offset = {offset}")
        idx = GraphIndex(quantize_u8_device(pts, slope, offset, device=device), None, "uint8", metric, max_degree=R, device=device)
    else:
        idx = GraphIndex(pts, None, dtype, metric, max_degree=R, device=device)
    idx.build(R=R, L=L, alpha=alpha, num_passes=2 if two_pass else 1)
    fileio.write_graph(index_dir, idx.graph())
    return idx


def load_index(metric, dtype, data_dir, index_dir, hnsw=False, device=0):
    """python/wrapper.py:96-116."""
    if hnsw:   # GraphIndex(data_path, index_path, is_hnsw=true) (graph_index.cpp:103-109): the v3 model file
        from .hnsw import load_hnsw_index
        return load_hnsw_index(metric, dtype, data_dir, index_dir, device=device)
    if metric not in _lib.METRIC:
        raise Exception("Invalid metric")
    if dtype not in _lib.DTYPE:
        raise Exception("Invalid data type")
    pts = fileio.read_bin(data_dir, dtype)
    if metric != "Euclidian" and fileio.np_dtype(dtype)().itemsize > 1:
        pts = normalize_rows_device(pts, device)   # graph_index.cpp:88-90: 4-byte non-metric points are normalised when the index loads
    return GraphIndex(pts, fileio.read_graph(index_dir), dtype, metric, device=device)
