"""CPU-only checks of the product's host side: the C-ABI library loads and exports every symbol
include/pann.h declares, its host logic (insertion order, batch schedule) equals the reference's, file
formats round-trip, and compute calls fail loudly without a GPU (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

import oracle_lib as O
from golden_util import GOLDEN

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from parlayann_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "pann.h")).read()
    declared = set(re.findall(r"\b(pann_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    L = _lib.lib()
    for s in declared:
        assert hasattr(L, s), s


def test_permutation_and_schedule_match_reference():
    import parlayann_b200 as P
    z = np.load(os.path.join(GOLDEN, "perms.npz"))
    for key in z.files:
        n = int(key.split("_")[1])
        assert np.array_equal(P.random_permutation(n), z[key]), n
    for n in (1, 5, 49, 50, 1000, 100000, 1000000, 60000000):
        assert P.batch_schedule(n) == O.batch_schedule(n), n
    s = P.batch_schedule(100000)
    assert len(s) == 60 and s[10] == (1023, 2047) and s[11] == (2047, 4047)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import parlayann_b200 as P
    pts = np.zeros((10, 16), np.uint8)
    with pytest.raises(P.PannError):
        P.GraphIndex(pts, None, "uint8", "Euclidian", max_degree=4)


def test_file_formats_roundtrip(tmp_path):
    from parlayann_b200 import fileio
    rng = np.random.default_rng(0)
    for dt, npdt in (("uint8", np.uint8), ("int8", np.int8), ("float", np.float32)):
        a = rng.integers(-100, 100, size=(37, 19)).astype(npdt)
        fileio.write_bin(tmp_path / "p.bin", a)
        assert np.array_equal(fileio.read_bin(tmp_path / "p.bin", dt), a)
    g = np.zeros((50, 9), np.uint32)
    g[:, 0] = rng.integers(0, 9, size=50)
    for i in range(50):
        g[i, 1:1 + g[i, 0]] = rng.integers(0, 50, size=g[i, 0])
    fileio.write_graph(tmp_path / "g", g)
    assert np.array_equal(fileio.read_graph(tmp_path / "g"), g)
    # header layout of graph.h:206-231: n, maxDeg, degrees, then edges
    raw = np.fromfile(tmp_path / "g", dtype=np.uint32)
    assert raw[0] == 50 and raw[1] == 8 and np.array_equal(raw[2:52], g[:, 0]) and raw.size == 52 + g[:, 0].sum()
    ids = rng.integers(0, 1000, size=(7, 5)).astype(np.int32)
    ds = rng.random((7, 5)).astype(np.float32)
    fileio.write_ibin(tmp_path / "gt", ids, ds)
    i2, d2 = fileio.read_ibin(tmp_path / "gt")
    assert np.array_equal(i2, ids) and np.array_equal(d2, ds)


def test_recall_with_ties():
    from parlayann_b200.harness import recall_with_ties
    gt_ids = np.array([[1, 2, 3, 4]], np.int32)
    gt_d = np.array([[0.0, 1.0, 1.0, 2.0]], np.float32)
    assert recall_with_ties(np.array([[1, 3]]), gt_ids, gt_d, 2) == 1.0   # 3 ties with the 2nd
    assert recall_with_ties(np.array([[1, 4]]), gt_ids, gt_d, 2) == 0.5


def test_random_starts_match_beamSearchRandom():
    """pann_random_starts = parlay::random_generator()[i] through std::uniform_int_distribution<long>(0, n-1)
    (beamSearch.h:327-332), pinned to the start points the reference's own beamSearchRandom used; the oracle with one start
    per query reproduces the reference's ids and counters."""
    from parlayann_b200.index import random_starts
    g = np.load(os.path.join(GOLDEN, "random_starts.npz"))
    n, d, nq, R, L, seed = (int(x) for x in g["params"])
    st = random_starts(nq, n)
    assert np.array_equal(st, g["starts"])
    base = O.gen_points("uint8", n, d, seed=seed)
    qs = O.gen_points("uint8", nq, d, seed=seed, stream=1)
    for si in range(3):
        k, beam = (int(x) for x in g[f"r{si}_kb"])
        r = O.search(base, g["graph"], qs, "uint8", k=k, beam=beam, starts=st.reshape(-1, 1))
        assert np.array_equal(r["ids"], g[f"r{si}_ids"])
        assert np.array_equal(r["visited"], g[f"r{si}_visited"]) and np.array_equal(r["cmps"], g[f"r{si}_distances"])


def test_ParlayANNpy_compat_surface():
    """parlayann_b200/compat/_ParlayANNpy.py exposes every name the reference's pybind module defines for this path
    (python/module.cpp:50-76, :113-126, :139-176) with the reference's argument names and order"""
    import importlib
    import inspect
    import sys
    sys.path.insert(0, os.path.join(ROOT, "parlayann_b200", "compat"))
    try:
        m = importlib.import_module("_ParlayANNpy")
        w = importlib.import_module("wrapper")
    finally:
        sys.path.pop(0)
    for t in ("float", "uint8", "int8"):
        for met in ("euclidian", "mips"):
            f = getattr(m, f"build_vamana_{t}_{met}_index")
            assert list(inspect.signature(f).parameters) == ["distance_metric", "data_file_path", "index_output_path", "graph_degree", "beam_width", "alpha", "two_pass"]
            f = getattr(m, f"build_hnsw_{t}_{met}_index")
            assert list(inspect.signature(f).parameters) == ["distance_metric", "data_file_path", "index_output_path", "graph_degree", "efc", "m_l", "alpha"]
            assert hasattr(m, f"build_hcnng_{t}_{met}_index") and hasattr(m, f"build_pynndescent_{t}_{met}_index")
    for cls in ("FloatEuclidianIndex", "FloatMipsIndex", "UInt8EuclidianIndex", "UInt8MipsIndex", "Int8EuclidianIndex", "Int8MipsIndex"):
        c = getattr(m, cls)
        assert list(inspect.signature(c.__init__).parameters)[1:] == ["index_path", "data_path", "hnsw"]
        assert list(inspect.signature(c.batch_search).parameters)[1:] == ["queries", "knn", "beam_width", "quant", "visit_limit"]
        assert list(inspect.signature(c.single_search).parameters)[1:] == ["q", "knn", "beam_width", "quant", "visit_limit"]
        assert list(inspect.signature(c.batch_search_from_string).parameters)[1:] == ["queries", "knn", "beam_width", "quant", "visit_limit"]
        assert list(inspect.signature(c.check_recall).parameters)[1:] == ["queries_file", "graph_file", "neighbors", "k"]
    assert (m.defaults.METRIC, m.defaults.ALPHA, m.defaults.GRAPH_DEGREE, m.defaults.BEAMWIDTH) == ("Euclidian", 1.2, 64, 128)
    for f in ("build_vamana_index", "build_hcnng_index", "build_pynndescent_index", "build_hnsw_index", "load_index"):
        assert callable(getattr(w, f))
    with pytest.raises(Exception, match="Invalid metric"):
        w.load_index("cosine", "float", "a", "b")
