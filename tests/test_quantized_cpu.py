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
