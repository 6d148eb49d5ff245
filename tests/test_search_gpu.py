"""K1 parity: the CUDA beam search (through the C ABI) against the CPU oracle on the same graph.
Every element type must match bit for bit: ids, distances, visited count, distance-comparison count
 — floats included, because the kernels add a row's terms in the reference's own order."""
import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


def _index(ds):
    import parlayann_b200 as P
    return P.GraphIndex(ds.base, ds.graph, ds.data_type, ds.metric)


def _check_exact(ds, idx, **qp):
    ids, dists, vis, cmps = idx.search(ds.queries, **qp)
    ref = O.search(ds.base, ds.graph, ds.queries, ds.data_type, ds.metric, **qp)
    assert np.array_equal(ids, ref["ids"]), f"ids differ for {qp}: {(ids != ref['ids']).any(axis=1).sum()} queries"
    assert np.array_equal(dists, ref["dists"]), f"dists differ for {qp}"
    assert np.array_equal(vis, ref["visited"]), f"visited counts differ for {qp}"
    assert np.array_equal(cmps, ref["cmps"]), f"distance-comparison counts differ for {qp}"


@pytest.mark.parametrize("beam", [10, 13, 20, 32, 33, 50, 64, 100, 128])
def test_u8_l2_beam_sweep(dataset, beam):
    ds = dataset("uint8")
    _check_exact(ds, _index(ds), k=10, beam=beam)


@pytest.mark.parametrize("beam", [150, 200, 300, 500])
def test_u8_l2_wide_beams_global_scratch(dataset, beam):
    ds = dataset("uint8")
    _check_exact(ds, _index(ds), k=10, beam=beam)


@pytest.mark.parametrize("beam,limit", [(10, 10), (15, 12), (20, 15), (40, 30), (100, 50), (10, 1), (64, 200)])
def test_u8_l2_visit_limits(dataset, beam, limit):
    # the reference sweep's visit-limit settings (check_nn_recall.h:243-254): limit < 2*beam disables the skip rule
    ds = dataset("uint8")
    _check_exact(ds, _index(ds), k=10, beam=beam, limit=limit)


@pytest.mark.parametrize("k,beam,cut", [(1, 10, 1.35), (10, 10, 1.35), (10, 100, 1.0), (10, 64, 1.1), (50, 100, 1.35), (100, 100, 1.35), (10, 1000, 10.0)])
def test_u8_l2_k_and_cut(dataset, k, beam, cut):
    ds = dataset("uint8")
    _check_exact(ds, _index(ds), k=k, beam=beam, cut=cut)


def test_u8_l2_degree_limit(dataset):
    ds = dataset("uint8")
    idx = _index(ds)
    for dl in (5, 16, 31, 32):
        _check_exact(ds, idx, k=10, beam=30, degree_limit=dl)
    # degree 1 walks a path and ends with fewer than k results: both sides refuse (beamSearch.h:416-419)
    import parlayann_b200 as P
    with pytest.raises(RuntimeError):
        O.search(ds.base, ds.graph, ds.queries, "uint8", k=10, beam=30, degree_limit=1)
    with pytest.raises(P.PannError):
        idx.search(ds.queries, k=10, beam=30, degree_limit=1)


@pytest.mark.parametrize("data_type,metric", [("int8", "Euclidian"), ("uint8", "mips"), ("int8", "mips")])
def test_other_integer_types_exact(dataset, data_type, metric):
    ds = dataset(data_type, metric, n=8000, nq=300)
    idx = _index(ds)
    for beam in (10, 40, 128):
        _check_exact(ds, idx, k=10, beam=beam)


@pytest.mark.parametrize("d", [100, 96, 200, 17, 2048])
def test_u8_odd_dimensions(dataset, d):
    # rows that are not a multiple of 16 bytes are zero padded on the device; d=2048 takes the generic (CPL=0) path
    ds = dataset("uint8", n=4000, d=d, nq=100)
    idx = _index(ds)
    _check_exact(ds, idx, k=10, beam=25)


def test_multi_start(dataset):
    # beam_search with several start points (HNSW layer search, beamSearch.h:217-223)
    ds = dataset("uint8")
    idx = _index(ds)
    _check_exact(ds, idx, k=10, beam=20, starts=(0, 17, 4242, 9))


def test_build_mode_visited_lists(dataset):
    # the build's search (vamana/index.h:249-259): k=0, beam=L, cut=0, limit=n, the query is a base row
    ds = dataset("uint8")
    idx = _index(ds)
    self_ids = np.arange(0, 2000, 7, dtype=np.uint32)
    got = idx.search_full(self_ids=self_ids, k=0, beam=64, cut=0.0, vcap=512)
    ref = O.search(ds.base, ds.graph, ds.base[self_ids], "uint8", k=0, beam=64, cut=0.0, selfs=self_ids, full=True, vcap=512)
    assert np.array_equal(got["nvis"], ref["nvis"]) and np.array_equal(got["nfr"], ref["nfr"])
    assert np.array_equal(got["cmps"], ref["cmps"])
    for i in range(len(self_ids)):
        nv, nf = ref["nvis"][i], ref["nfr"][i]
        assert np.array_equal(got["vis_ids"][i, :nv], ref["vis_ids"][i, :nv])
        assert np.array_equal(got["vis_dists"][i, :nv], ref["vis_dists"][i, :nv])
        assert np.array_equal(got["fr_ids"][i, :nf], ref["fr_ids"][i, :nf])


@pytest.mark.parametrize("metric,d", [("Euclidian", 128), ("mips", 200), ("Euclidian", 96)])
def test_float_exact(dataset, metric, d):
    # K1 accumulates one row per lane in the reference's own dimension order with unfused multiply/add, so float
    # searches are bit-identical to the CPU's too (north_star asks for 1e-5 relative; this is stricter)
    ds = dataset("float", metric, n=8000, d=d, nq=300)
    idx = _index(ds)
    for beam in (10, 40, 100):
        _check_exact(ds, idx, k=10, beam=beam)


def test_errors(dataset):
    import parlayann_b200 as P
    ds = dataset("uint8")
    idx = _index(ds)
    with pytest.raises(P.PannError):  # k > beam (beamSearch.h:549-553)
        idx.search(ds.queries, k=20, beam=10)
    with pytest.raises(P.PannError):  # no start point (beamSearch.h:38-41)
        idx.search(ds.queries, k=10, beam=10, starts=())
    with pytest.raises(P.PannError):  # fewer than k results (beamSearch.h:416-419): limit 0 visits nothing
        idx.search(ds.queries, k=10, beam=10, limit=0)
    ids, _, vis, _ = idx.search(ds.queries[:0], k=10, beam=10)  # empty batch
    assert ids.shape == (0, 10)


def test_resident_queries_match_host_path(dataset):
    ds = dataset("uint8")
    idx = _index(ds)
    idx.upload_queries(ds.queries)
    idx.search_resident(k=10, beam=13)
    a = idx.fetch_results()
    b = idx.search(ds.queries, k=10, beam=13)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    ms, launches = idx.last_timing()
    assert launches >= 1 and ms > 0


def test_synthetic_device_generator_matches_cpu(dataset):
    import parlayann_b200 as P
    for dt, kind, d, r, nc in (("uint8", "uint8", 128, 24, 256), ("int8", "int8", 64, 24, 256), ("float", "float", 128, 24, 256),
                               ("float", "float_norm", 96, 24, 256), ("float", "float", 200, 32, 0)):
        idx = P.GraphIndex(data_type=dt, max_degree=8, synthetic=dict(n=3000, d=d, kind=kind, seed=99, r=r, nc=nc))
        cpu = O.gen_points(kind, 3000, d, seed=99, r=r, nc=nc)
        assert np.array_equal(idx.points(), cpu), (dt, kind, d)
        idx.synthetic_queries(200, kind=kind, seed=99, stream=1, r=r, nc=nc)
        assert np.array_equal(idx.download_queries(), O.gen_points(kind, 200, d, seed=99, stream=1, r=r, nc=nc))


def test_search_matches_reference_golden():
    # the CUDA path against outputs of the UNMODIFIED reference (tests/golden, made by oracle/_ref/ref_driver)
    import parlayann_b200 as P
    from golden_util import CASES, SEARCHES, golden
    for name in CASES:
        g = golden(name)
        idx = P.GraphIndex(g.base, g.graph, g.data_type, g.metric)
        ran = 0
        for si, s in enumerate(SEARCHES):
            if f"s{si}_ids" not in g.z:
                continue
            ids, _, vis, cmps = idx.search(g.queries, **s)
            assert np.array_equal(ids, g.z[f"s{si}_ids"]), (name, s)
            assert np.array_equal(vis, g.z[f"s{si}_visited"]) and np.array_equal(cmps, g.z[f"s{si}_distances"]), (name, s)
            ran += 1
        assert ran >= 4
        # build-mode visited lists
        cnt = g.z["self_counts"]
        got = idx.search_full(self_ids=np.arange(g.nq, dtype=np.uint32), k=0, beam=g.L, cut=0.0, vcap=2048)
        assert np.array_equal(got["nvis"], cnt[:, 1]) and np.array_equal(got["cmps"], cnt[:, 2])
        vi = g.split_lists(cnt[:, 1], g.z["self_visited"])
        for i in range(g.nq):
            assert np.array_equal(got["vis_ids"][i, :cnt[i, 1]], vi[i][:, 0])
            assert np.array_equal(got["vis_dists"][i, :cnt[i, 1]].view(np.uint32), vi[i][:, 1])


def test_pipelined_host_batch_identical(dataset, monkeypatch):
    # large pinned host batches go through the chunk pipeline of pann_search_batch (copies under the neighbouring
    # chunks' kernels, two compute streams): same results as the plain path, and as the oracle on the original rows
    torch = pytest.importorskip("torch")
    ds = dataset("uint8")
    idx = _index(ds)
    reps = 300000 // len(ds.queries) + 1
    nq = reps * len(ds.queries)                       # > 2 x 128K: the pipeline engages; ragged last chunk
    hq = torch.from_numpy(np.tile(ds.queries, (reps, 1))).pin_memory()

    def pinned(shape, dt):
        return torch.empty(shape, dtype=dt).pin_memory().numpy()

    def run():
        out = (pinned((nq, 10), torch.int32).view(np.uint32), pinned((nq, 10), torch.float32),
               pinned(nq, torch.int32).view(np.uint32), pinned(nq, torch.int32).view(np.uint32))
        for o in out:
            o[...] = 0
        idx.search(hq.numpy(), k=10, beam=16, out=out)
        return out, idx.last_timing()[1]
    piped, nl = run()
    assert nl >= 3, "the chunk pipeline did not engage"
    monkeypatch.setenv("PANN_NO_PIPELINE", "1")
    plain, nl0 = run()
    assert nl0 == 1
    for a, b in zip(piped, plain):
        assert np.array_equal(a, b)
    ref = O.search(ds.base, ds.graph, ds.queries, ds.data_type, ds.metric, k=10, beam=16)
    n0 = len(ds.queries)
    for block in (0, reps - 1):
        sl = slice(block * n0, (block + 1) * n0)
        assert np.array_equal(piped[0][sl], ref["ids"]) and np.array_equal(piped[1][sl], ref["dists"])
        assert np.array_equal(piped[2][sl], ref["visited"]) and np.array_equal(piped[3][sl], ref["cmps"])


@pytest.mark.parametrize("beam,env", [(100, None), (128, None), (40, "PANN_FORCE_GLOBAL_SCRATCH")])
def test_pipelined_wide_beam_identical(dataset, monkeypatch, beam, env):
    # wide beams keep the hash filter and the visited list in a global slab: the two compute streams of the chunk pipeline
    # must not share one (two chunk kernels are resident at the same time) — pipelined == plain == oracle
    torch = pytest.importorskip("torch")
    ds = dataset("uint8")
    idx = _index(ds)
    if env:
        monkeypatch.setenv(env, "1")
    reps = 270000 // len(ds.queries) + 1
    nq = reps * len(ds.queries)
    hq = torch.from_numpy(np.tile(ds.queries, (reps, 1))).pin_memory()

    def pinned(shape, dt):
        return torch.empty(shape, dtype=dt).pin_memory().numpy()

    def run():
        out = (pinned((nq, 10), torch.int32).view(np.uint32), pinned((nq, 10), torch.float32),
               pinned(nq, torch.int32).view(np.uint32), pinned(nq, torch.int32).view(np.uint32))
        for o in out:
            o[...] = 0
        idx.search(hq.numpy(), k=10, beam=beam, out=out)
        return out, idx.last_timing()[1]
    piped, nl = run()
    assert nl >= 3, "the chunk pipeline did not engage"
    monkeypatch.setenv("PANN_NO_PIPELINE", "1")
    plain, _ = run()
    for a, b in zip(piped, plain):
        assert np.array_equal(a, b)
    ref = O.search(ds.base, ds.graph, ds.queries, ds.data_type, ds.metric, k=10, beam=beam)
    n0 = len(ds.queries)
    for block in (0, reps // 2, reps - 1):
        sl = slice(block * n0, (block + 1) * n0)
        assert np.array_equal(piped[0][sl], ref["ids"]) and np.array_equal(piped[2][sl], ref["visited"]) and np.array_equal(piped[3][sl], ref["cmps"])


# ---- one start list per query (beamSearchRandom, beamSearch.h:309-351; HNSW's per-query entry points)
def _check_exact_starts(ds, idx, starts, nstarts_q=None, **qp):
    ids, dists, vis, cmps = idx.search(ds.queries, starts=starts, nstarts_q=nstarts_q, **qp)
    ref = O.search(ds.base, ds.graph, ds.queries, ds.data_type, ds.metric, starts=starts, nstarts_q=nstarts_q, **qp)
    assert np.array_equal(ids, ref["ids"]) and np.array_equal(dists, ref["dists"])
    assert np.array_equal(vis, ref["visited"]) and np.array_equal(cmps, ref["cmps"])


@pytest.mark.parametrize("beam", [10, 24, 32, 40, 100])
def test_per_query_single_start(dataset, beam):
    import parlayann_b200 as P
    from parlayann_b200.index import random_starts
    ds = dataset("uint8")
    idx = _index(ds)
    st = random_starts(len(ds.queries), ds.n).reshape(-1, 1)
    assert st.max() < ds.n and len(np.unique(st)) > len(st) // 2
    _check_exact_starts(ds, idx, st, k=10, beam=beam)


@pytest.mark.parametrize("dt,metric", [("int8", "mips"), ("float", "Euclidian")])
def test_per_query_ragged_start_lists(dataset, dt, metric):
    ds = dataset(dt, metric, n=8000, d=100 if dt == "float" else 128, nq=300)
    idx = _index(ds)
    rng = np.random.default_rng(5)
    st = rng.integers(0, ds.n, size=(len(ds.queries), 6), dtype=np.uint32)
    cnt = rng.integers(1, 7, size=len(ds.queries), dtype=np.uint32)
    for beam in (12, 48):   # K1s and the general kernel
        ids, dists, vis, cmps = idx.search(ds.queries, starts=st, nstarts_q=cnt, k=10, beam=beam)
        ref = O.search(ds.base, ds.graph, ds.queries, ds.data_type, ds.metric, starts=st, nstarts_q=cnt, k=10, beam=beam)
        assert np.array_equal(vis, ref["visited"]) and np.array_equal(cmps, ref["cmps"])
        if dt == "float":
            assert np.allclose(dists, ref["dists"], rtol=1e-5, atol=0)
        else:
            assert np.array_equal(ids, ref["ids"]) and np.array_equal(dists, ref["dists"])


@pytest.mark.parametrize("beam,k", [(1, 1), (1, 0), (4, 4), (3, 0)])
def test_more_starts_than_beam(dataset, beam, k):
    # HNSW's ef = 1 descent from several entrances (HNSW.hpp:869-881): the frontier starts longer than the beam and the
    # first merge truncates it; whole frontier and visited list against the oracle
    ds = dataset("uint8")
    idx = _index(ds)
    rng = np.random.default_rng(9)
    st = rng.integers(0, ds.n, size=(len(ds.queries), 7), dtype=np.uint32)
    cnt = rng.integers(1, 8, size=len(ds.queries), dtype=np.uint32)
    got = idx.search_full(ds.queries, k=k, beam=beam, starts=st, nstarts_q=cnt, vcap=256)
    ref = O.search(ds.base, ds.graph, ds.queries, "uint8", k=k, beam=beam, starts=st, nstarts_q=cnt, full=True, fcap=7, vcap=256)
    assert np.array_equal(got["nvis"], ref["nvis"]) and np.array_equal(got["nfr"], ref["nfr"]) and np.array_equal(got["cmps"], ref["cmps"])
    for i in range(len(ds.queries)):
        nv, nf = ref["nvis"][i], ref["nfr"][i]
        assert np.array_equal(got["vis_ids"][i, :nv], ref["vis_ids"][i, :nv])
        assert np.array_equal(got["fr_ids"][i, :nf], ref["fr_ids"][i, :nf]) and np.array_equal(got["fr_dists"][i, :nf], ref["fr_dists"][i, :nf])
    # without the visited list the narrow-beam kernel takes the same searches
    ids, dists, vis, cmps = idx.search(ds.queries, starts=st, nstarts_q=cnt, k=k if k else 1, beam=beam) if k else (None,) * 4
    if k:
        assert np.array_equal(ids, ref["fr_ids"][:, :k]) and np.array_equal(vis, ref["nvis"]) and np.array_equal(cmps, ref["cmps"])


def test_more_starts_than_beam_with_cut_rule_is_refused(dataset):
    import parlayann_b200 as P
    ds = dataset("uint8")
    idx = _index(ds)
    st = np.tile(np.arange(12, dtype=np.uint32), (len(ds.queries), 1))
    with pytest.raises(P.PannError):
        idx.search(ds.queries, starts=st, k=2, beam=10)


def test_beamSearchRandom_golden():
    # the reference's own beamSearchRandom (ref_driver random -> tests/golden/random_starts.npz): same start points, ids, counters
    import os
    import parlayann_b200 as P
    from parlayann_b200.index import random_starts
    from golden_util import GOLDEN
    g = np.load(os.path.join(GOLDEN, "random_starts.npz"))
    n, d, nq, R, L, seed = (int(x) for x in g["params"])
    base = O.gen_points("uint8", n, d, seed=seed)
    qs = O.gen_points("uint8", nq, d, seed=seed, stream=1)
    idx = P.GraphIndex(base, g["graph"], "uint8", "Euclidian")
    st = random_starts(nq, n)
    assert np.array_equal(st, g["starts"])
    for si in range(3):
        k, beam = (int(x) for x in g[f"r{si}_kb"])
        ids, _, vis, cmps = idx.search(qs, k=k, beam=beam, starts=st.reshape(-1, 1))
        assert np.array_equal(ids, g[f"r{si}_ids"])
        assert np.array_equal(vis, g[f"r{si}_visited"]) and np.array_equal(cmps, g[f"r{si}_distances"])


@pytest.mark.parametrize("data_type,metric", [("uint8", "Euclidian"), ("int8", "mips")])
def test_small_batch_wide_variant_identical(dataset, monkeypatch, data_type, metric):
    """K1s WIDE (both 32-neighbour rounds of a hop filtered first, then all row gathers in flight together — the kernel small
    batches get) against the default kernel and the oracle: ids, distances, visited and comparison counters"""
    ds = dataset(data_type, metric, n=20000, d=128, nq=300, R=64, L=96)
    idx = _index(ds)
    for kw in (dict(k=10, beam=24), dict(k=10, beam=12, limit=20), dict(k=5, beam=32, cut=1.1), dict(k=10, beam=16, degree_limit=40)):
        monkeypatch.setenv("PANN_K1S_WIDE", "0")
        a = idx.search(ds.queries, **kw)
        monkeypatch.setenv("PANN_K1S_WIDE", "1")
        b = idx.search(ds.queries, **kw)
        ref = O.search(ds.base, ds.graph, ds.queries, data_type, metric, **kw)
        for x, y in zip(a, b):
            assert np.array_equal(x, y), kw
        assert np.array_equal(b[0], ref["ids"]) and np.array_equal(b[2], ref["visited"]) and np.array_equal(b[3], ref["cmps"]), kw
