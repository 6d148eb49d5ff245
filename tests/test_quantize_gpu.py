"""csrc/quantize.cu — the reference's point conversions (normalise, 1-byte scalar quantisers) as kernels behind the C ABI.
Same bytes as the oracle's C++ restatement and as the numpy mirror that tests/test_quantized_cpu.py pins to the reference's
own constructors (utils/euclidian_point.h:183-237, utils/mips_point.h:115-124, :421-462)."""
import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind,n,d", [("float", 5000, 100), ("float_norm", 3000, 96), ("float", 257, 7)])
def test_u8_quantiser_equals_oracle(kind, n, d):
    from parlayann_b200 import index as I
    base = O.gen_points(kind, n, d, seed=31)
    base[::7] *= 3.5                      # wider range, negative values stay
    slope, offset = I.quant_params_u8_device(base)
    s2, o2 = O.quant_params_u8(base)
    assert np.float32(slope) == np.float32(s2) and offset == o2
    assert I.quant_params_u8(base) == (slope, offset)
    got = I.quantize_u8_device(base, slope, offset)
    assert np.array_equal(got, O.quantize_u8(base, slope, offset)) and np.array_equal(got, I.quantize_u8(base, slope, offset))
    for s, o in ((np.float32(0.37), -3), (np.float32(1.0), 1), (np.float32(200.0), 17)):
        assert np.array_equal(I.quantize_u8_device(base, s, o), O.quantize_u8(base, s, o))
    # halves: x * slope exactly on .5 must round away from zero
    h = (np.arange(-40, 40, dtype=np.float32) + np.float32(0.5)).reshape(8, 10)
    assert np.array_equal(I.quantize_u8_device(h, np.float32(1.0), -20), I.quantize_u8(h, np.float32(1.0), -20))
    # all-integer non-negative data below 256 maps unscaled (slope 1, offset 0)
    ints = O.gen_points("uint8", 500, 64, seed=4).astype(np.float32)
    assert I.quant_params_u8_device(ints) == (np.float32(1.0), 0) == I.quant_params_u8(ints)
    assert np.array_equal(I.quantize_u8_device(ints, np.float32(1.0), 0), ints.astype(np.uint8))


@pytest.mark.parametrize("n,d", [(4000, 200), (1000, 33)])
def test_mips_normalise_and_int8_quantiser_equal_oracle(n, d):
    from parlayann_b200 import index as I
    base = O.gen_points("float", n, d, seed=32)
    base[5] = 0                            # a zero row keeps norm 1
    nb = I.normalize_rows_device(base)
    assert np.array_equal(nb.view(np.uint32), O.normalize_rows(base).view(np.uint32))
    assert np.array_equal(nb.view(np.uint32), I.normalize_rows(base).view(np.uint32))
    for trim in (True, False):
        mv = I.quant_params_mips8_device(nb, trim=trim)
        assert np.float32(mv) == np.float32(I.quant_params_mips8(nb, trim=trim))
    mv = I.quant_params_mips8_device(nb, trim=True)
    assert np.float32(mv) == np.float32(O.quant_params_mips8(nb))
    got = I.quantize_mips8_device(nb, mv)
    assert np.array_equal(got, O.quantize_mips8(nb, mv)) and np.array_equal(got, I.quantize_mips8(nb, mv))
    # strided input rows (a column slice) and the error paths
    wide = np.zeros((n, d + 5), np.float32); wide[:, :d] = nb
    assert np.array_equal(I.quantize_mips8_device(wide[:, :d], mv), got)
    from parlayann_b200 import _lib
    import ctypes as C
    out = np.zeros((n, d + 3), np.int8)          # strided rows straight through the C ABI, input and output
    assert _lib.lib().pann_rows_quantize_mips8(0, wide.ctypes.data_as(C.c_void_p), C.c_uint64(n), C.c_uint32(d), C.c_uint64(wide.strides[0]),
                                               C.c_float(float(mv)), out.ctypes.data_as(C.c_void_p), C.c_uint64(out.strides[0])) == 0
    assert np.array_equal(out[:, :d], got) and not out[:, d:].any()
    assert _lib.lib().pann_rows_normalize(0, None, C.c_uint64(3), C.c_uint32(4), C.c_uint64(16)) != 0
    assert _lib.lib().pann_rows_normalize(0, nb.ctypes.data_as(C.c_void_p), C.c_uint64(3), C.c_uint32(d), C.c_uint64(4)) != 0


def test_device_quantisers_match_the_reference_fixtures():
    """the committed fixtures written by the reference's own quantising constructors (tests/golden/quant_*.npz, the same
    checks tests/test_quantized_cpu.py makes of the oracle and the numpy mirror)"""
    import zlib
    import test_quantized_cpu as T
    from parlayann_b200 import index as I
    for name in sorted(T.QCASES):
        z, base, qs = T.load(name)
        slope, offset = I.quant_params_u8_device(base)
        assert slope == z["slope"] and offset == int(z["offset"])
        qb = I.quantize_u8_device(base, slope, offset)
        assert np.array_equal(qb[:64], z["qbase_head"]) and np.uint32(zlib.crc32(qb.tobytes())) == z["qbase_crc"]
        assert np.array_equal(I.quantize_u8_device(qs, slope, offset), z["qquery"])
    for name in sorted(T.MIPS_QCASES):
        z, base, qs = T.load_mips(name)
        nb, nq = I.normalize_rows_device(base), I.normalize_rows_device(qs)
        assert np.array_equal(nb[:64], z["nbase_head"]) and np.uint32(zlib.crc32(nq.tobytes())) == z["nquery_crc"]
        mv = I.quant_params_mips8_device(nb)
        assert mv == z["max_val"]
        qb = I.quantize_mips8_device(nb, mv)
        assert np.array_equal(qb[:64], z["qbase_head"]) and np.uint32(zlib.crc32(qb.tobytes())) == z["qbase_crc"]
        assert np.array_equal(I.quantize_mips8_device(nq, mv), z["qquery"])
