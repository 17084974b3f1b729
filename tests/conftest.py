"""Test configuration.

`-m "not gpu"`: oracle vs golden vectors, host logic, C-ABI symbol exports (runs without a GPU).
`-m gpu`      : parity of the CUDA path against the oracle through the C ABI (needs a B200).

Only tests/ (and smoke()/bench.py's CPU-baseline legs) may load anything under oracle/.
"""
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from whisper_rust_b200 import api, synth  # noqa: E402

ORACLE_LIB = ROOT / "oracle" / "_ref" / "libwhisper_ref.so"
GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


_PROBE_SIGS = {
    "ref_probe_mel": (C.POINTER(C.c_float), [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "ref_probe_embd_enc": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "ref_probe_embd_conv": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "ref_probe_kv_cross_k": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_int64]),
    "ref_probe_kv_cross_v": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_int64]),
    "ref_probe_counters": (None, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "ref_probe_aheads_qks": (C.POINTER(C.c_float), [C.c_void_p, C.POINTER(C.c_int64)]),
}

_B200_SIGS = {
    "whisper_b200_full_batch": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, api.whisper_full_params,
                                          C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int)]),
    "whisper_b200_encode_batch": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.c_int]),
    "whisper_b200_decode_batch": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int32), C.POINTER(C.c_int), C.c_int]),
    "whisper_b200_get_mel": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "whisper_b200_get_embd_enc": (C.c_int64, [C.c_void_p, C.c_int, C.c_void_p, C.c_int64]),
    "whisper_b200_get_kv_cross": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_int64]),
    "whisper_b200_get_aheads_qk": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_int64]),
    "whisper_b200_get_counters": (None, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "whisper_b200_kernel_launches": (C.c_uint64, []),
    "whisper_b200_device_sync": (None, [C.c_void_p]),
    "whisper_b200_test_gemm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.c_int, C.c_void_p, C.c_int]),
    "whisper_b200_test_gemm_ex": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.c_int, C.c_void_p, C.c_int, C.c_int]),
    "whisper_b200_bench_gemm": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "whisper_b200_test_enc_attn": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]),
    "whisper_b200_bench_enc_attn": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "whisper_b200_test_mel": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "whisper_b200_test_cross_attn": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
}


def declare(lib, sigs):
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


@pytest.fixture(scope="session")
def ref_lib():
    """The oracle: the reference's own CPU engine compiled from /root/reference by oracle/Makefile."""
    if not ORACLE_LIB.exists():
        pytest.skip(f"{ORACLE_LIB} not built (make -C oracle)")
    lib = api.load_library(ORACLE_LIB)
    declare(lib, _PROBE_SIGS)
    api.install_quiet_logging(lib)
    return lib


@pytest.fixture(scope="session")
def prod_lib():
    """The product library.  Must exist: there is no fallback of any kind."""
    lib = api.load_library()
    declare(lib, _B200_SIGS)
    if not os.environ.get("WB200_TEST_LOG"):   # WB200_TEST_LOG=1: leave the library's stderr logging on (debugging a failure)
        api.install_quiet_logging(lib)
    return lib


_MODEL_CACHE = {}


def model_blob(arch: str, recipe: synth.InitRecipe = synth.InitRecipe()) -> bytes:
    key = (arch, recipe)
    if key not in _MODEL_CACHE:
        _MODEL_CACHE[key] = synth.model_bytes(arch, recipe)
    return _MODEL_CACHE[key]


@pytest.fixture(scope="session")
def blobs():
    return model_blob


GGML_TYPE_NAMES = {2: "q4_0", 3: "q4_1", 6: "q5_0", 7: "q5_1", 8: "q8_0", 10: "q2_K", 11: "q3_K", 12: "q4_K", 13: "q5_K", 14: "q6_K"}


def ggml_codec(lib):
    """(quantizer, dequantizer) backed by the reference's own ggml_quantize_chunk / dequantize_row_* (ggml-quants.c) in the oracle
    library: the block codec synth.write_model takes for the formats it does not pack itself."""
    lib.ggml_cpu_init()                                                     # the f16 -> f32 table dequantize_row_* reads
    lib.ggml_quantize_chunk.restype = C.c_size_t
    lib.ggml_quantize_chunk.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
    lib.ggml_row_size.restype = C.c_size_t
    lib.ggml_row_size.argtypes = [C.c_int, C.c_int64]

    def quantizer(x, ttype):
        x = np.ascontiguousarray(x, dtype=np.float32)
        cols = x.shape[-1]
        rows = x.size // cols
        buf = np.zeros(lib.ggml_row_size(ttype, cols) * rows, dtype=np.uint8)
        assert lib.ggml_quantize_chunk(ttype, x.ctypes.data, buf.ctypes.data, 0, rows, cols, None) == buf.size
        return buf.tobytes()

    def dequantizer(raw, ttype, n):
        fn = getattr(lib, "dequantize_row_" + GGML_TYPE_NAMES[ttype])
        fn.restype, fn.argtypes = None, [C.c_void_p, C.c_void_p, C.c_int64]
        src = np.frombuffer(raw, dtype=np.uint8)
        out = np.empty(n, dtype=np.float32)
        fn(src.ctypes.data, out.ctypes.data, n)
        return out

    return quantizer, dequantizer


def ref_mel(lib, st):
    n_len, n_org, n_mel = C.c_int(), C.c_int(), C.c_int()
    p = lib.ref_probe_mel(st.ptr, C.byref(n_len), C.byref(n_org), C.byref(n_mel))
    a = np.ctypeslib.as_array(p, shape=(n_mel.value, n_len.value)).copy()
    return a, n_org.value


def prod_mel(lib, st):
    n_len, n_org, n_mel = C.c_int(), C.c_int(), C.c_int()
    n = lib.whisper_b200_get_mel(st.ptr, None, 0, C.byref(n_len), C.byref(n_org), C.byref(n_mel))
    a = np.empty((n_mel.value, n_len.value), dtype=np.float32)
    assert lib.whisper_b200_get_mel(st.ptr, a.ctypes.data, a.size, None, None, None) == n
    return a, n_org.value


def ref_embd_enc(lib, st):
    ne = (C.c_int64 * 4)()
    nb = lib.ref_probe_embd_enc(st.ptr, None, 0, ne)
    a = np.empty((ne[1], ne[0]), dtype=np.float32)
    assert nb == a.nbytes
    lib.ref_probe_embd_enc(st.ptr, a.ctypes.data, a.nbytes, ne)
    return a


def prod_embd_enc(lib, ctx, window=0):
    n = lib.whisper_b200_get_embd_enc(ctx.ptr, window, None, 0)
    S = ctx.lib.whisper_model_n_audio_state(ctx.ptr)
    a = np.empty((n // S, S), dtype=np.float32)
    assert lib.whisper_b200_get_embd_enc(ctx.ptr, window, a.ctypes.data, a.size) == n
    return a


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))
