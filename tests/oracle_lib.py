"""ctypes access to the CHECKERS: oracle/_build/liboracle.so (our C++ restatement) and
oracle/_ref/ref_driver (the unmodified reference, when it was built).  Test infrastructure only."""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
DT = {"float": 0, "uint8": 1, "int8": 2}
NPDT = {"float": np.float32, "uint8": np.uint8, "int8": np.int8}
METRIC = {"Euclidian": 0, "mips": 1}
GEN = {"uint8": 0, "int8": 1, "float": 2, "float_norm": 3}

_lib = None


def lib():
    global _lib
    if _lib is None:
        p = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
        if not os.path.exists(p):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "oracle"])
        _lib = C.CDLL(p)
        _lib.oracle_distance.restype = C.c_float
        _lib.oracle_batch_schedule.restype = C.c_int64
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def nthreads():
    return lib().oracle_num_threads()


def gen_points(kind, n, d, seed=1234, stream=0, r=24, nc=256, row0=0):
    out = np.empty((n, d), dtype=np.float32 if kind.startswith("float") else NPDT[kind])
    lib().oracle_gen_points(GEN[kind], C.c_uint64(seed), C.c_uint64(stream), C.c_uint64(row0), C.c_uint64(n),
                            C.c_uint32(d), C.c_uint32(r), C.c_uint32(nc), _p(out), C.c_uint64(out.strides[0]),
                            nthreads())
    return out


def random_permutation(n):
    out = np.empty(n, dtype=np.int32)
    lib().oracle_random_permutation(C.c_int64(n), _p(out))
    return out


def batch_schedule(n, m=None):
    m = n if m is None else m
    lo = np.zeros(4096, dtype=np.int64)
    hi = np.zeros(4096, dtype=np.int64)
    nb = lib().oracle_batch_schedule(C.c_uint64(n), C.c_uint64(m), _p(lo), _p(hi), C.c_int64(4096))
    return [(int(lo[i]), int(hi[i])) for i in range(nb)]


def search(points, graph, queries, data_type, metric="Euclidian", k=10, beam=10, cut=1.35, limit=None,
           degree_limit=None, starts=(0,), selfs=None, full=False, fcap=None, vcap=4096, nstarts_q=None):
    """starts: one list for the batch, or a 2-D [nq, stride] array of per-query lists (nstarts_q: how many of each row count)"""
    n, d = points.shape
    nq = queries.shape[0]
    maxdeg = graph.shape[1] - 1
    limit = n if limit is None or limit < 0 else limit
    degree_limit = maxdeg if degree_limit is None or degree_limit < 0 else degree_limit
    st = np.ascontiguousarray(starts, dtype=np.uint32)
    sstride = st.shape[1] if st.ndim == 2 else 0
    nsq = None if nstarts_q is None else np.ascontiguousarray(nstarts_q, dtype=np.uint32)
    ids = np.zeros((nq, k), dtype=np.uint32)
    dists = np.zeros((nq, k), dtype=np.float32)
    vis = np.zeros(nq, dtype=np.uint32)
    cmps = np.zeros(nq, dtype=np.uint32)
    fcap = beam if fcap is None else fcap
    fr_i = fr_d = nfr = vi_i = vi_d = nvi = None
    if full:
        fr_i = np.zeros((nq, fcap), dtype=np.uint32); fr_d = np.zeros((nq, fcap), dtype=np.float32)
        vi_i = np.zeros((nq, vcap), dtype=np.uint32); vi_d = np.zeros((nq, vcap), dtype=np.float32)
        nfr = np.zeros(nq, dtype=np.uint32); nvi = np.zeros(nq, dtype=np.uint32)
    sf = None if selfs is None else np.ascontiguousarray(selfs, dtype=np.int64)
    points = np.ascontiguousarray(points); queries = np.ascontiguousarray(queries)
    graph = np.ascontiguousarray(graph, dtype=np.uint32)
    rc = lib().oracle_search_batch_starts(DT[data_type], METRIC[metric], _p(points), C.c_uint64(n), C.c_uint32(d),
                                   C.c_uint64(points.strides[0]), _p(graph), C.c_uint32(maxdeg), _p(queries),
                                   C.c_uint64(nq), C.c_uint64(queries.strides[0]), _p(sf), _p(st), C.c_uint32(sstride or len(st)),
                                   C.c_uint32(sstride), _p(nsq),
                                   C.c_long(k), C.c_long(beam), C.c_double(cut), C.c_long(limit), C.c_long(degree_limit),
                                   _p(ids), _p(dists), _p(vis), _p(cmps),
                                   C.c_uint32(fcap), _p(fr_i), _p(fr_d), _p(nfr),
                                   C.c_uint32(vcap), _p(vi_i), _p(vi_d), _p(nvi), nthreads())
    if rc:
        raise RuntimeError("oracle: a query returned fewer than k results (the reference aborts here)")
    out = dict(ids=ids, dists=dists, visited=vis, cmps=cmps)
    if full:
        out.update(fr_ids=fr_i, fr_dists=fr_d, nfr=nfr, vis_ids=vi_i, vis_dists=vi_d, nvis=nvi)
    return out


def vamana_build(points, data_type, metric="Euclidian", R=32, L=64, alpha=1.2, passes=1, sort_neighbors=True):
    n, d = points.shape
    points = np.ascontiguousarray(points)
    g = np.zeros((n, R + 1), dtype=np.uint32)
    sv = np.zeros(n, dtype=np.uint32)
    sd = np.zeros(n, dtype=np.uint32)
    lib().oracle_vamana_build(DT[data_type], METRIC[metric], _p(points), C.c_uint64(n), C.c_uint32(d),
                              C.c_uint64(points.strides[0]), C.c_uint32(R), C.c_uint32(L), C.c_double(alpha),
                              passes, int(sort_neighbors), _p(g), _p(sv), _p(sd), nthreads())
    return g, sv, sd


def vamana_batch_insert(points, graph, inserts, data_type, metric="Euclidian", L=64, alpha=1.2, random_order=False, max_fraction=0.02):
    """knn_index::batch_insert of `inserts` into `graph` (modified in place); returns the stats increments (visited, distances)"""
    n, d = points.shape
    points = np.ascontiguousarray(points)
    assert graph.dtype == np.uint32 and graph.flags["C_CONTIGUOUS"]
    ins = np.ascontiguousarray(inserts, dtype=np.uint32)
    sv = np.zeros(n, dtype=np.uint32); sd = np.zeros(n, dtype=np.uint32)
    lib().oracle_vamana_batch_insert(DT[data_type], METRIC[metric], _p(points), C.c_uint64(n), C.c_uint32(d), C.c_uint64(points.strides[0]),
                                     C.c_uint32(graph.shape[1] - 1), C.c_uint32(L), C.c_double(alpha), _p(ins), C.c_uint64(len(ins)),
                                     int(bool(random_order)), C.c_double(max_fraction), _p(graph), _p(sv), _p(sd), nthreads())
    return sv, sd


def groundtruth(base, queries, data_type, metric="Euclidian", k=100):
    base = np.ascontiguousarray(base); queries = np.ascontiguousarray(queries)
    nq = queries.shape[0]
    ids = np.zeros((nq, k), dtype=np.int32)
    dists = np.zeros((nq, k), dtype=np.float32)
    rc = lib().oracle_groundtruth(DT[data_type], METRIC[metric], _p(base), C.c_uint64(base.shape[0]), _p(queries),
                                  C.c_uint64(nq), C.c_uint32(base.shape[1]), C.c_uint64(base.strides[0]),
                                  C.c_uint64(queries.strides[0]), C.c_uint32(k), _p(ids), _p(dists), nthreads())
    if rc:
        raise RuntimeError("oracle_groundtruth: k > nb")
    return ids, dists


# ---- the real reference (only where oracle/_ref/ref_driver exists and runs on this CPU) ----
_ref_path = []


def ref_driver_path():
    """the first of ref_driver (-march=native of the build container), ref_driver_v4 (x86-64-v4), ref_driver_v3 that runs on
    this CPU (a binary built for an ISA extension this CPU lacks dies with SIGILL in its self test)"""
    if _ref_path:
        return _ref_path[0]
    for name in ("ref_driver", "ref_driver_v4", "ref_driver_v3"):
        p = os.path.join(ORACLE_DIR, "_ref", name)
        if os.path.exists(p):
            try:
                if subprocess.run([p, "selftest"], capture_output=True, timeout=60).returncode == 0:
                    _ref_path.append(p)
                    return p
            except Exception:
                pass
    _ref_path.append(None)
    return None


def ref_driver_info():
    """which reference binary ref_run uses and the ISA level it was compiled for (oracle/_ref/BUILD_INFO)"""
    p = ref_driver_path()
    if p is None:
        return "unavailable"
    name = os.path.basename(p)
    try:
        for line in open(os.path.join(ORACLE_DIR, "_ref", "BUILD_INFO")):
            if line.startswith(name + ":"):
                return line.strip()
    except OSError:
        pass
    return name


def ref_run(args, threads=None):
    p = ref_driver_path()
    if p is None:
        raise RuntimeError("oracle/_ref/ref_driver is not available")
    env = dict(os.environ)
    if threads:
        env["PARLAY_NUM_THREADS"] = str(threads)
    r = subprocess.run([p] + [str(a) for a in args], capture_output=True, text=True, env=env)
    if r.returncode != 0:
        raise RuntimeError(f"ref_driver failed ({r.returncode}): {r.stdout[-2000:]} {r.stderr[-2000:]}")
    return r.stdout


def quant_params_u8(points):
    """Euclidian_Point<uint8_t>::generate_parameters on float rows -> (slope float32, offset int32)"""
    points = np.ascontiguousarray(points, dtype=np.float32)
    slope = C.c_float(0); offset = C.c_int32(0)
    lib().oracle_quant_params_u8(_p(points), C.c_uint64(points.shape[0]), C.c_uint32(points.shape[1]), C.c_uint64(points.strides[0]),
                                 C.byref(slope), C.byref(offset))
    return np.float32(slope.value), int(offset.value)


def quantize_u8(points, slope, offset):
    points = np.ascontiguousarray(points, dtype=np.float32)
    out = np.empty(points.shape, dtype=np.uint8)
    lib().oracle_quantize_u8(_p(points), C.c_uint64(points.shape[0]), C.c_uint32(points.shape[1]), C.c_uint64(points.strides[0]),
                             C.c_float(float(slope)), C.c_int32(offset), _p(out))
    return out


def normalize_rows(points):
    """Mips_Point<float>::normalize over every row (mips_point.h:115-124) -> a new float32 array"""
    out = np.array(points, dtype=np.float32, order="C")
    lib().oracle_normalize_f32(_p(out), C.c_uint64(out.shape[0]), C.c_uint32(out.shape[1]), C.c_uint64(out.strides[0]))
    return out


def quant_params_mips8(points, trim=True):
    """Quantized_Mips_Point<8,trim>::generate_parameters -> max_val float32"""
    points = np.ascontiguousarray(points, dtype=np.float32)
    f = lib().oracle_quant_params_mips8
    f.restype = C.c_float
    return np.float32(f(_p(points), C.c_uint64(points.shape[0]), C.c_uint32(points.shape[1]), C.c_uint64(points.strides[0]), C.c_int(int(trim))))


def quantize_mips8(points, max_val):
    points = np.ascontiguousarray(points, dtype=np.float32)
    out = np.empty(points.shape, dtype=np.int8)
    lib().oracle_quantize_mips8(_p(points), C.c_uint64(points.shape[0]), C.c_uint32(points.shape[1]), C.c_uint64(points.strides[0]),
                                C.c_float(float(max_val)), _p(out))
    return out

def search_rerank(points, qpoints, graph, queries, qqueries, data_type="float", metric="Euclidian", k=10, beam=10, cut=1.35,
                  limit=None, degree_limit=None, rerank_factor=100, qdata_type="uint8"):
    """beam_search_rerank with a 1-byte quantised copy (beamSearch.h:390-454): search qpoints/qqueries (uint8), then the
    exact re-rank of the first k*rerank_factor frontier entries against points/queries."""
    r = search(qpoints, graph, qqueries, qdata_type, metric, k=k, beam=beam, cut=cut, limit=limit, degree_limit=degree_limit, full=True)
    points = np.ascontiguousarray(points); queries = np.ascontiguousarray(queries)
    nq = queries.shape[0]
    ids = np.zeros((nq, k), dtype=np.uint32); dists = np.zeros((nq, k), dtype=np.float32)
    fcap = r["fr_ids"].shape[1]
    rc = lib().oracle_rerank(DT[data_type], METRIC[metric], _p(points), C.c_uint64(points.shape[0]), C.c_uint32(points.shape[1]),
                             C.c_uint64(points.strides[0]), _p(queries), C.c_uint64(nq), C.c_uint64(queries.strides[0]),
                             _p(r["fr_ids"]), _p(r["nfr"]), C.c_uint32(fcap), C.c_long(k), C.c_long(rerank_factor), _p(ids), _p(dists), nthreads())
    if rc:
        raise RuntimeError("oracle: a query returned fewer than k results (the reference aborts here)")
    return dict(ids=ids, dists=dists, visited=r["visited"], cmps=r["cmps"])


# ---- HNSW (algorithms/HNSW/HNSW.hpp): our C++ restatement, and the reference itself (oracle/_ref/ref_hnsw)
def hnsw_build(points, data_type, metric, m, efc, m_l=0.36, alpha=1.0, levels=None):
    """-> parlayann_b200.hnsw_file.HnswModel built by the oracle (levels of one worker unless given)"""
    from parlayann_b200.hnsw_file import HnswModel
    points = np.ascontiguousarray(points)
    n, d = points.shape
    sizes = np.zeros(3, np.uint64)
    lv = None if levels is None else np.ascontiguousarray(levels, dtype=np.uint32)
    lib().oracle_hnsw_build(DT[data_type], METRIC[metric], _p(points), C.c_uint64(n), C.c_uint32(d), C.c_uint64(points.strides[0]),
                            C.c_float(m_l), C.c_uint32(m), C.c_uint32(efc), C.c_float(alpha), _p(lv), _p(sizes), nthreads())
    level = np.zeros(n, np.uint32); counts = np.zeros(int(sizes[0]), np.uint64); edges = np.zeros(int(sizes[1]), np.uint32)
    ent = np.zeros(int(sizes[2]), np.uint32)
    lib().oracle_hnsw_export(_p(level), _p(counts), _p(edges), _p(ent))
    return HnswModel(d, m_l, m, efc, alpha, level, counts, edges, ent)


def hnsw_levels(n, m_l):
    out = np.zeros(n, np.uint32)
    lib().oracle_hnsw_levels(C.c_uint64(n), C.c_float(m_l), _p(out))
    return out


def hnsw_search(model, points, queries, data_type, metric, k=10, ef=50, limit=-1):
    """HNSW::search over `model` -> dict(ids [nq, ef], dists, cnt, cmps); cnt = entries the reference returns (k + ties)"""
    M = model
    lib().oracle_hnsw_import(C.c_uint32(M.n), C.c_uint32(M.dim), C.c_float(M.m_l), C.c_uint32(M.m), C.c_uint32(M.efc), C.c_float(M.alpha),
                             _p(M.level), _p(M.counts), _p(M.edges), _p(M.entrance), C.c_uint64(len(M.entrance)))
    points = np.ascontiguousarray(points); queries = np.ascontiguousarray(queries)
    nq = queries.shape[0]
    ids = np.zeros((nq, ef), np.uint32); dists = np.zeros((nq, ef), np.float32)
    cnt = np.zeros(nq, np.uint32); cmps = np.zeros(nq, np.uint32)
    lib().oracle_hnsw_search(DT[data_type], METRIC[metric], _p(points), C.c_uint64(points.strides[0]), _p(queries), C.c_uint64(nq),
                             C.c_uint64(queries.strides[0]), C.c_uint32(k), C.c_uint32(ef), C.c_long(limit), C.c_uint32(ef),
                             _p(ids), _p(dists), _p(cnt), _p(cmps), nthreads())
    return dict(ids=ids, dists=dists, cnt=cnt, cmps=cmps)


def ref_hnsw_path():
    p = os.path.join(ORACLE_DIR, "_ref", "ref_hnsw")
    try:
        if os.path.exists(p) and subprocess.run([p, "selftest"], capture_output=True, timeout=60).returncode == 0:
            return p
    except Exception:
        pass
    return None


def ref_hnsw_run(args, threads=None):
    p = ref_hnsw_path()
    if p is None:
        raise RuntimeError("oracle/_ref/ref_hnsw is not available")
    env = dict(os.environ)
    if threads:
        env["PARLAY_NUM_THREADS"] = str(threads)
    r = subprocess.run([p] + [str(a) for a in args], capture_output=True, text=True, env=env)
    if r.returncode != 0:
        raise RuntimeError(f"ref_hnsw failed ({r.returncode}): {r.stdout[-2000:]} {r.stderr[-2000:]}")
    return r.stdout


def parse_hnsw_results(path, nq):
    """ref_hnsw search output: per query u32 count, then (id, dist bits) pairs"""
    a = np.fromfile(path, dtype=np.uint32)
    out, pos = [], 0
    for _ in range(nq):
        c = int(a[pos]); pos += 1
        pr = a[pos:pos + 2 * c].reshape(c, 2); pos += 2 * c
        out.append((pr[:, 0].copy(), pr[:, 1].copy().view(np.float32)))
    assert pos == a.size
    return out
