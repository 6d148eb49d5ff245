"""The oracle's 1-byte scalar quantiser and search + exact re-rank (beam_search_rerank with differing byte widths,
beamSearch.h:390-454; Euclidian_Point<uint8_t>, euclidian_point.h:182-237) against outputs of the UNMODIFIED reference
(tests/golden/quant_*.npz, made by tests/golden/make_golden_quant.py through ref_driver qsearch).  Also pins the
host-side quantiser of the Python binding (parlayann_b200.index.quantize_u8), which must produce the same bytes."""
import os
import zlib

import numpy as np
import pytest

import oracle_lib as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
QCASES = {"quant_f32_l2": (8192, 128, 200), "quant_f32_l2_d100": (4096, 100, 100)}
BEAMS = [10, 16, 40, 100]


def load(name):
    n, d, nq = QCASES[name]
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    base = O.gen_points("float", n, d, seed=77)
    qs = O.gen_points("float", nq, d, seed=77, stream=1)
    return z, base, qs


@pytest.mark.parametrize("name", sorted(QCASES))
def test_quantiser_matches_reference(name):
    z, base, qs = load(name)
    slope, offset = O.quant_params_u8(base)
    assert slope == z["slope"] and offset == int(z["offset"])
    qb = O.quantize_u8(base, slope, offset)
    assert np.array_equal(qb[:64], z["qbase_head"]) and np.uint32(zlib.crc32(qb.tobytes())) == z["qbase_crc"]
    assert np.array_equal(O.quantize_u8(qs, slope, offset), z["qquery"])
    # the binding's numpy quantiser (host logic of batch_search(quant=True)) produces the same bytes
    from parlayann_b200.index import quant_params_u8, quantize_u8
    s2, o2 = quant_params_u8(base)
    assert s2 == slope and o2 == offset
    assert np.array_equal(quantize_u8(base, s2, o2), qb) and np.array_equal(quantize_u8(qs, s2, o2), z["qquery"])


@pytest.mark.parametrize("name", sorted(QCASES))
def test_search_rerank_matches_reference(name):
    z, base, qs = load(name)
    slope, offset = O.quant_params_u8(base)
    qb, qq = O.quantize_u8(base, slope, offset), O.quantize_u8(qs, slope, offset)
    for beam in BEAMS:
        r = O.search_rerank(base, qb, z["graph"], qs, qq, k=10, beam=beam)
        assert np.array_equal(r["ids"], z[f"ids_{beam}"]), f"beam {beam}"
        assert np.array_equal(r["visited"], z[f"visited_{beam}"]) and np.array_equal(r["cmps"], z[f"cmps_{beam}"])


def test_quantiser_edge_cases():
    from parlayann_b200.index import quant_params_u8, quantize_u8
    # all-integer non-negative data below 256 is taken as is (slope 1, offset 0: euclidian_point.h:229-233)
    a = np.arange(0, 200, dtype=np.float32).reshape(10, 20)
    for qp, qz in ((O.quant_params_u8, O.quantize_u8), (quant_params_u8, quantize_u8)):
        s, o = qp(a)
        assert s == 1.0 and o == 0
        assert np.array_equal(qz(a, s, o), a.astype(np.uint8))
    # halves round away from zero (std::round), values beyond the fitted range clamp to [0, 255]
    b = np.array([[-1.0, 0.0, 0.5, 1.5, 2.5, 254.5, 300.0, -7.0]], dtype=np.float32)
    assert np.array_equal(quantize_u8(b, np.float32(1.0), 1), O.quantize_u8(b, np.float32(1.0), 1))
    assert np.array_equal(quantize_u8(b, np.float32(0.37), -3), O.quantize_u8(b, np.float32(0.37), -3))


# ---- float MIPS data: Mips_Point<float>::normalize + Quantized_Mips_Point<8,true> (mips_point.h:115-124, :288-462)
MIPS_QCASES = {"quant_f32_mips": (8192, 128, 200), "quant_f32_mips_d200": (4096, 200, 100)}


def load_mips(name):
    n, d, nq = MIPS_QCASES[name]
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return z, O.gen_points("float", n, d, seed=78), O.gen_points("float", nq, d, seed=78, stream=1)


@pytest.mark.parametrize("name", sorted(MIPS_QCASES))
def test_mips_quantiser_matches_reference(name):
    from parlayann_b200.index import normalize_rows, quant_params_mips8, quantize_mips8
    z, base, qs = load_mips(name)
    for norm, qpar, quant in ((O.normalize_rows, O.quant_params_mips8, O.quantize_mips8), (normalize_rows, quant_params_mips8, quantize_mips8)):
        nb, nq = norm(base), norm(qs)
        assert np.array_equal(nb[:64], z["nbase_head"]) and np.uint32(zlib.crc32(nq.tobytes())) == z["nquery_crc"]
        mv = qpar(nb)
        assert mv == z["max_val"]
        qb = quant(nb, mv)
        assert qb.dtype == np.int8 and np.array_equal(qb[:64], z["qbase_head"]) and np.uint32(zlib.crc32(qb.tobytes())) == z["qbase_crc"]
        assert np.array_equal(quant(nq, mv), z["qquery"])


@pytest.mark.parametrize("name", sorted(MIPS_QCASES))
def test_mips_build_and_search_rerank_match_reference(name):
    """Quantized_Mips_Point<8>'s distance is the int8 inner product (mips_point.h:339-346), so the reference's build and
    search over that range equal the int8 MIPS oracle over the quantised bytes: graph (ref_driver qbuild, alpha overridden
    as builder.cpp:50-53 does) and re-ranked results (ref_driver qsearch) are identical."""
    z, base, qs = load_mips(name)
    nb, nq = O.normalize_rows(base), O.normalize_rows(qs)
    mv = O.quant_params_mips8(nb)
    qb, qq = O.quantize_mips8(nb, mv), O.quantize_mips8(nq, mv)
    alpha = float(np.float32(1.0 if base.shape[1] < 100 else .98))
    g = O.vamana_build(qb, "int8", "mips", R=32, L=64, alpha=alpha, passes=1)[0]
    live = np.arange(g.shape[1])[None, :] <= g[:, :1]
    assert np.array_equal(g[:, 0], z["graph"][:, 0]) and np.array_equal(g[live], z["graph"][live])
    for beam in BEAMS:
        r = O.search_rerank(nb, qb, z["graph"], nq, qq, data_type="float", metric="mips", k=10, beam=beam, qdata_type="int8")
        assert np.array_equal(r["ids"], z[f"ids_{beam}"]), f"beam {beam}"
        assert np.array_equal(r["visited"], z[f"visited_{beam}"]) and np.array_equal(r["cmps"], z[f"cmps_{beam}"])


def test_mips_quantiser_edge_cases():
    from parlayann_b200.index import normalize_rows, quant_params_mips8, quantize_mips8
    a = np.array([[0, 0, 0, 0], [3, 4, 0, 0], [1e-20, 0, 0, 0], [-2, 2, -2, 2]], dtype=np.float32)   # zero row keeps norm 1
    assert np.array_equal(normalize_rows(a), O.normalize_rows(a))
    b = np.array([[-1.0, -0.5, -0.1968504, 0.0, 0.0039370, 0.1968504, 0.5, 0.7]], dtype=np.float32)  # caps and halves
    for mv in (np.float32(0.5), np.float32(0.3171855)):
        assert np.array_equal(quantize_mips8(b, mv), O.quantize_mips8(b, mv))
    for trim in (True, False):
        assert quant_params_mips8(b, trim) == O.quant_params_mips8(b, trim)
    rng = np.random.default_rng(3)
    c = rng.standard_normal((300, 37)).astype(np.float32)
    assert quant_params_mips8(c) == O.quant_params_mips8(c) and quant_params_mips8(c, False) == O.quant_params_mips8(c, False)
