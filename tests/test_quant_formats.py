"""The model loader's expansion of quantised ggml tensors (csrc/model.cpp: dequantize_blocks / expand_k_block) against the reference's
dequantize_row_* (ggml/src/ggml-quants.c:260-420 legacy blocks, 706-1720 K super-blocks).  Host logic: runs without a GPU.
The bar is bit-identical f32 output, on blocks the reference's own quantiser produced and on random bit patterns."""
import ctypes as C

import numpy as np
import pytest

from conftest import GGML_TYPE_NAMES, ggml_codec

# bytes per block, elements per block (ggml-common.h)
BLOCK = {2: (18, 32), 3: (20, 32), 6: (22, 32), 7: (24, 32), 8: (34, 32), 10: (84, 256), 11: (110, 256), 12: (144, 256), 13: (176, 256),
         14: (210, 256)}
# byte offsets of the f16 fields inside a block (kept finite in the random-bytes test)
HALF_FIELDS = {2: (0,), 3: (0, 2), 6: (0,), 7: (0, 2), 8: (0,), 10: (80, 82), 11: (108,), 12: (0, 2), 13: (0, 2), 14: (208,)}


@pytest.fixture(scope="module")
def expand(prod_lib):
    prod_lib.whisper_b200_test_dequantize.restype = C.c_int
    prod_lib.whisper_b200_test_dequantize.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_void_p]

    def run(ttype, raw, n):
        src = np.frombuffer(raw, dtype=np.uint8)
        out = np.full(n, np.nan, dtype=np.float32)
        assert prod_lib.whisper_b200_test_dequantize(ttype, src.ctypes.data, n, out.ctypes.data) == 0
        return out

    return run


@pytest.mark.parametrize("ttype", sorted(BLOCK))
def test_expansion_of_reference_quantised_rows_is_bit_identical(ref_lib, expand, ttype):
    quantizer, dequantizer = ggml_codec(ref_lib)
    rng = np.random.default_rng(100 + ttype)
    rows, cols = 48, 1280
    x = (rng.standard_normal((rows, cols)) * np.exp(2.0 * rng.standard_normal((rows, 1)))).astype(np.float32)   # row scales over ~4 decades
    x[3] = 0.0                  # an all-zero row (d = 0)
    x[5, :256] = 0.0            # an all-zero block inside a live row
    x[7] = np.abs(x[7])         # one-signed rows: the minimum / offset fields at their extremes
    x[9] = -np.abs(x[9])
    raw = quantizer(x, ttype)
    assert len(raw) == rows * cols // BLOCK[ttype][1] * BLOCK[ttype][0]
    want = dequantizer(raw, ttype, x.size)
    got = expand(ttype, raw, x.size)
    np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32), err_msg=GGML_TYPE_NAMES[ttype])
    assert np.abs(want - x.ravel()).max() < 0.4 * np.abs(x).max()        # and the codec round trip is a quantisation, not garbage


@pytest.mark.parametrize("ttype", sorted(BLOCK))
def test_expansion_of_random_bit_patterns_is_bit_identical(ref_lib, expand, ttype):
    """every packed field at arbitrary values (scale tables with all high bits, every level), which a quantiser's output does not reach"""
    _, dequantizer = ggml_codec(ref_lib)
    rng = np.random.default_rng(200 + ttype)
    bs, ne = BLOCK[ttype]
    nb = 512
    blocks = rng.integers(0, 256, size=(nb, bs), dtype=np.uint8)
    for off in HALF_FIELDS[ttype]:                                       # finite, signed, normal and subnormal f16 scales
        h = np.clip(rng.standard_normal(nb) * np.exp(3.0 * rng.standard_normal(nb)), -6.0e4, 6.0e4).astype(np.float16)
        h[::17] = np.float16(6e-8) * rng.integers(1, 100, size=h[::17].shape)
        blocks[:, off:off + 2] = h.view(np.uint8).reshape(nb, 2)
    raw = blocks.tobytes()
    want = dequantizer(raw, ttype, nb * ne)
    got = expand(ttype, raw, nb * ne)
    assert np.isfinite(want).all()
    np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32), err_msg=GGML_TYPE_NAMES[ttype])


def test_expansion_refuses_other_types_and_partial_blocks(prod_lib, expand):
    buf = np.zeros(4096, dtype=np.uint8)
    out = np.zeros(4096, dtype=np.float32)
    f = prod_lib.whisper_b200_test_dequantize
    for ttype in (0, 1, 4, 5, 9, 15, 16, 20, 30):                        # f32 / f16 are not block types; 16+ are the IQ codebook types
        assert f(ttype, buf.ctypes.data, 256, out.ctypes.data) == 1
    assert f(12, buf.ctypes.data, 128, out.ctypes.data) == 1             # half a super-block
    assert f(8, buf.ctypes.data, 48, out.ctypes.data) == 1
    assert f(8, None, 32, out.ctypes.data) == 1
