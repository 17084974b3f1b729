"""ctypes access to the VAD part of whisper.h (wh: whisper_vad_*; the crate's safe layer has no binding for it, so the tests drive the C
functions directly) and to the default-state entry points whisper_full / whisper_full_parallel, the only callers that read
whisper_full_params.vad (whisper.cpp:7720, 7748)."""
import ctypes as C
import os
import tempfile

import numpy as np

from whisper_rust_b200 import api, synth


class vad_context_params(C.Structure):
    _fields_ = [("n_threads", C.c_int), ("use_gpu", C.c_bool), ("gpu_device", C.c_int)]


def declare(lib):
    V, I = C.c_void_p, C.c_int
    sigs = {
        "whisper_vad_default_context_params": (vad_context_params, []),
        "whisper_vad_default_params": (api.whisper_vad_params, []),
        "whisper_vad_init_from_file_with_params": (V, [C.c_char_p, vad_context_params]),
        "whisper_vad_detect_speech": (C.c_bool, [V, V, I]),
        "whisper_vad_n_probs": (I, [V]),
        "whisper_vad_probs": (C.POINTER(C.c_float), [V]),
        "whisper_vad_segments_from_probs": (V, [V, api.whisper_vad_params]),
        "whisper_vad_segments_from_samples": (V, [V, api.whisper_vad_params, V, I]),
        "whisper_vad_segments_n_segments": (I, [V]),
        "whisper_vad_segments_get_segment_t0": (C.c_float, [V, I]),
        "whisper_vad_segments_get_segment_t1": (C.c_float, [V, I]),
        "whisper_vad_free_segments": (None, [V]),
        "whisper_vad_free": (None, [V]),
        "whisper_init_from_buffer_with_params": (V, [V, C.c_size_t, api.whisper_context_params]),
        "whisper_full": (I, [V, api.whisper_full_params, C.POINTER(C.c_float), I]),
        "whisper_full_parallel": (I, [V, api.whisper_full_params, C.POINTER(C.c_float), I, I]),
        "whisper_full_n_segments": (I, [V]),
        "whisper_full_get_segment_t0": (C.c_int64, [V, I]),
        "whisper_full_get_segment_t1": (C.c_int64, [V, I]),
        "whisper_full_get_segment_text": (C.c_char_p, [V, I]),
        "whisper_full_n_tokens": (I, [V, I]),
        "whisper_full_get_token_id": (C.c_int32, [V, I, I]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


_MODEL_PATH = {}


def vad_model_path(**kw):
    """the synthetic Silero-layout checkpoint as a file (the API takes a path: whisper_full_params.vad_model_path)"""
    key = tuple(sorted(kw.items()))
    if key not in _MODEL_PATH:
        fd, p = tempfile.mkstemp(prefix="wb200_vad_", suffix=".bin")
        with os.fdopen(fd, "wb") as f:
            synth.write_vad_model(f, **kw)
        _MODEL_PATH[key] = p
    return _MODEL_PATH[key]


class Vad:
    def __init__(self, lib, gpu_device=0, **kw):
        self.lib = declare(lib)
        p = lib.whisper_vad_default_context_params()
        p.gpu_device = gpu_device
        self.ptr = lib.whisper_vad_init_from_file_with_params(vad_model_path(**kw).encode(), p)
        assert self.ptr, "VAD model failed to load"

    def probs(self, pcm):
        a = np.ascontiguousarray(pcm, dtype=np.float32)
        assert self.lib.whisper_vad_detect_speech(self.ptr, a.ctypes.data, len(a))
        n = self.lib.whisper_vad_n_probs(self.ptr)
        return np.ctypeslib.as_array(self.lib.whisper_vad_probs(self.ptr), shape=(n,)).copy()

    def segments(self, params=None):
        params = params if params is not None else self.lib.whisper_vad_default_params()
        s = self.lib.whisper_vad_segments_from_probs(self.ptr, params)
        assert s
        out = [(self.lib.whisper_vad_segments_get_segment_t0(s, i), self.lib.whisper_vad_segments_get_segment_t1(s, i))
               for i in range(self.lib.whisper_vad_segments_n_segments(s))]
        self.lib.whisper_vad_free_segments(s)
        return out

    def free(self):
        self.lib.whisper_vad_free(self.ptr)
        self.ptr = None


def default_state_context(lib, blob, use_gpu):
    """whisper_init_from_buffer_with_params: a context WITH its default state (what whisper_full needs)"""
    declare(lib)
    cp = lib.whisper_context_default_params()
    cp.use_gpu = use_gpu
    buf = (C.c_char * len(blob)).from_buffer_copy(blob)
    ptr = lib.whisper_init_from_buffer_with_params(C.cast(buf, C.c_void_p), len(blob), cp)
    assert ptr
    return ptr, buf


def default_state_segments(lib, ctx):
    out = []
    for i in range(lib.whisper_full_n_segments(ctx)):
        out.append((lib.whisper_full_get_segment_t0(ctx, i), lib.whisper_full_get_segment_t1(ctx, i), lib.whisper_full_get_segment_text(ctx, i),
                    [lib.whisper_full_get_token_id(ctx, i, j) for j in range(lib.whisper_full_n_tokens(ctx, i))]))
    return out
