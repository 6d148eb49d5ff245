"""K1 over a 1-byte quantised copy + K7 exact re-rank (pann_search_rerank_batch) against the UNMODIFIED reference's
outputs (tests/golden/quant_*.npz: qsearchAll with quantised ranges, i.e. -quantize_mode 1) and the oracle."""
import numpy as np
import pytest

import oracle_lib as O
from test_quantized_cpu import BEAMS, QCASES, load

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
    ref_graph = O.vamana_build(qb, "uint8", "Euclidian", R=16, L=32, alpha=1.2, passes=1, sort_neighbors=True)[0]
    g = fileio.read_graph(gp)
    live = np.arange(g.shape[1])[None, :] <= g[:, :1]          # slot 0 = degree; the file format drops the slots beyond it
    assert np.array_equal(g[:, 0], ref_graph[:, 0]) and np.array_equal(g[live], ref_graph[live])
    idx = P.load_index("Euclidian", "float", bp, gp)
    ids, dists = idx.batch_search(qs, 10, 30, quant=True, visit_limit=len(base))
    want = O.search_rerank(base, qb, g, qs, quantize_u8(qs, *quant_params_u8(base)), k=10, beam=30)
    assert np.array_equal(ids, want["ids"]) and np.array_equal(dists, want["dists"])
    with pytest.raises(NotImplementedError):
        P.build_vamana_index("mips", "float", bp, gp, 16, 32, 1.0, False)
