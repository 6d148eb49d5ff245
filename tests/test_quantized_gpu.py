"""K1 over a 1-byte quantised copy + K7 exact re-rank (pann_search_rerank_batch) against the UNMODIFIED reference's
outputs (tests/golden/quant_*.npz: qsearchAll with quantised ranges, i.e. -quantize_mode 1) and the oracle."""
import os

import numpy as np
import pytest

import oracle_lib as O
from test_quantized_cpu import BEAMS, GOLDEN, QCASES, load

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", sorted(QCASES))
def test_search_rerank_matches_reference(name):
    import parlayann_b200 as P
    from parlayann_b200.index import quant_params_u8, quantize_u8
    z, base, qs = load(name)
    slope, offset = quant_params_u8(base)
    qb, qq = quantize_u8(base, slope, offset), quantize_u8(qs, slope, offset)
    idx = P.GraphIndex(base, z["graph"], "float", "Euclidian")
    qidx = P.GraphIndex(qb, z["graph"], "uint8", "Euclidian")
    for beam in BEAMS:
        ids, dists, vis, cmps = idx.search_rerank(qidx, qs, qq, k=10, beam=beam)
        assert np.array_equal(ids, z[f"ids_{beam}"]), f"beam {beam}: {(ids != z[f'ids_{beam}']).any(axis=1).sum()} queries differ"
        assert np.array_equal(vis, z[f"visited_{beam}"]) and np.array_equal(cmps, z[f"cmps_{beam}"])
        ref = O.search_rerank(base, qb, z["graph"], qs, qq, k=10, beam=beam)
        assert np.array_equal(dists, ref["dists"])          # exact float distances, summed in the reference's order
    # rerank_factor limits how much of the frontier is re-ranked (beamSearch.h:428)
    ids, dists, _, _ = idx.search_rerank(qidx, qs, qq, k=5, beam=40, rerank_factor=2)
    ref = O.search_rerank(base, qb, z["graph"], qs, qq, k=5, beam=40, rerank_factor=2)
    assert np.array_equal(ids, ref["ids"]) and np.array_equal(dists, ref["dists"])
    # the reference's Python surface: batch_search(quant=True) builds the 1-byte copy itself (graph_index.cpp:141-186)
    ids2, dists2 = idx.batch_search(qs, 10, 40, quant=True, visit_limit=len(base))
    assert np.array_equal(ids2, z["ids_40"])


def test_rerank_errors():
    import parlayann_b200 as P
    z, base, qs = load("quant_f32_l2_d100")
    idx = P.GraphIndex(base, z["graph"], "float", "Euclidian")
    other = P.GraphIndex(base[:100].astype(np.uint8), None, "uint8", "Euclidian", max_degree=32)
    with pytest.raises(P.PannError):   # the two point sets must describe the same points
        idx.search_rerank(other, qs, qs.astype(np.uint8), k=10, beam=10)


def test_builder_quantises_float_data(tmp_path):
    # python/builder.cpp:60-72: 4-byte Euclidean data is built on its 1-byte scalar-quantised copy; the graph must be the
    # oracle's build over exactly those bytes, and the saved index serves quant=True searches
    import parlayann_b200 as P
    from parlayann_b200 import fileio
    from parlayann_b200.index import quant_params_u8, quantize_u8
    base = O.gen_points("float", 3000, 64, seed=5)
    qs = O.gen_points("float", 50, 64, seed=5, stream=1)
    bp, gp = str(tmp_path / "b.fbin"), str(tmp_path / "g.graph")
    fileio.write_bin(bp, base)
    P.build_vamana_index("Euclidian", "float", bp, gp, 16, 32, 1.2, False)
    qb = quantize_u8(base, *quant_params_u8(base))
    ref_graph = O.vamana_build(qb, "uint8", "Euclidian", R=16, L=32, alpha=float(np.float32(1.2)), passes=1, sort_neighbors=True)[0]  # builder.cpp: float alpha
    g = fileio.read_graph(gp)
    live = np.arange(g.shape[1])[None, :] <= g[:, :1]          # slot 0 = degree; the file format drops the slots beyond it
    assert np.array_equal(g[:, 0], ref_graph[:, 0]) and np.array_equal(g[live], ref_graph[live])
    idx = P.load_index("Euclidian", "float", bp, gp)
    ids, dists = idx.batch_search(qs, 10, 30, quant=True, visit_limit=len(base))
    want = O.search_rerank(base, qb, g, qs, quantize_u8(qs, *quant_params_u8(base)), k=10, beam=30)
    assert np.array_equal(ids, want["ids"]) and np.array_equal(dists, want["dists"])
    with pytest.raises(NotImplementedError):    # 1-byte MIPS rows: the reference normalises them in their integer type
        P.build_vamana_index("mips", "uint8", bp, gp, 16, 32, 1.0, False)


# ---- float MIPS data: normalised rows, Quantized_Mips_Point<8,true> copy (graph_index.cpp:88-95, :166-181; builder.cpp:45-54, :73-86)
MIPS_QCASES = {"quant_f32_mips": (8192, 128, 200), "quant_f32_mips_d200": (4096, 200, 100)}


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(MIPS_QCASES))
def test_mips_quantised_search_matches_reference(name, tmp_path):
    """load_index + batch_search(quant=True) on float MIPS data against the UNMODIFIED reference's outputs
    (tests/golden/quant_f32_mips*.npz: ref_driver qsearch = qsearchAll over Quantized_Mips_Point<8,true> + exact re-rank)."""
    import parlayann_b200 as P
    from parlayann_b200 import fileio
    n, d, nq = MIPS_QCASES[name]
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    base = O.gen_points("float", n, d, seed=78)
    qs = O.gen_points("float", nq, d, seed=78, stream=1)
    bp, gp = str(tmp_path / "b.fbin"), str(tmp_path / "g.graph")
    fileio.write_bin(bp, base); fileio.write_graph(gp, z["graph"])
    idx = P.load_index("mips", "float", bp, gp)
    nb, nqs = O.normalize_rows(base), O.normalize_rows(qs)
    assert np.array_equal(idx.points(), nb)                      # normalised at load time
    mv = O.quant_params_mips8(nb)
    qb, qq = O.quantize_mips8(nb, mv), O.quantize_mips8(nqs, mv)
    for beam in (10, 16, 40, 100):
        ids, dists = idx.batch_search(qs, 10, beam, quant=True, visit_limit=n)
        assert np.array_equal(ids, z[f"ids_{beam}"]), f"beam {beam}"
        want = O.search_rerank(nb, qb, z["graph"], nqs, qq, data_type="float", metric="mips", k=10, beam=beam, qdata_type="int8")
        assert np.array_equal(ids, want["ids"]) and np.allclose(dists, want["dists"], rtol=1e-5, atol=0)
    # quant=False searches the normalised float rows with the caller's queries as they are (graph_index.cpp:184)
    ids, dists = idx.batch_search(qs, 10, 40, quant=False, visit_limit=n)
    want = O.search(nb, z["graph"], qs, "float", "mips", k=10, beam=40)
    assert np.array_equal(ids, want["ids"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(MIPS_QCASES))
def test_mips_build_vamana_index_matches_reference(name, tmp_path):
    """build_vamana_index("mips", "float") = normalise, override alpha, build over the Quantized_Mips_Point<8,true> copy:
    the graph file equals the one the reference's builder logic produces (ref_driver qbuild golden)."""
    import parlayann_b200 as P
    from parlayann_b200 import fileio
    n, d, _ = MIPS_QCASES[name]
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    bp, gp = str(tmp_path / "b.fbin"), str(tmp_path / "g.graph")
    fileio.write_bin(bp, O.gen_points("float", n, d, seed=78))
    P.build_vamana_index("mips", "float", bp, gp, 32, 64, 1.2, False)
    g, ref = fileio.read_graph(gp), z["graph"]
    live = np.arange(g.shape[1])[None, :] <= g[:, :1]
    assert np.array_equal(g[:, 0], ref[:, 0]) and np.array_equal(g[live], ref[live])
