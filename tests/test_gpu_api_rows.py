"""SURVEY §8(a) rows that sit beside the main loop, product vs oracle through the same C entry points:
F6 tokenizer, F4 language auto-detection, B1 set_mel, and whisper_full_parallel (§8(f) item 1)."""
import ctypes as C

import numpy as np
import pytest

from conftest import model_blob, prod_mel
from test_gpu_stages import RICH, _same
from whisper_rust_b200 import FullParams, SamplingStrategy, WhisperContext, WhisperContextParameters, api, synth

pytestmark = pytest.mark.gpu


def _ctx(lib, gpu):
    return WhisperContext.new_from_buffer_with_params(model_blob("mini", RICH), WhisperContextParameters(use_gpu=gpu), lib=lib)


def test_tokenizer_equals_oracle(prod_lib, ref_lib):
    """whisper_tokenize (wc:3288-3336, 3973-3986): GPT-2 pre-split + greedy longest match; -(needed) when the buffer is short"""
    a, b = _ctx(ref_lib, False), _ctx(prod_lib, True)
    for text in ["hello world", " t123 t456, t7!", "it's  12345 we'll they've I'm", "héllo 世界", "", "   ", "a\nb\tc  d", "x" * 300]:
        assert a.tokenize(text, 512) == b.tokenize(text, 512), text
    for lib, ctx in ((ref_lib, a), (prod_lib, b)):
        buf = (C.c_int32 * 4)()
        assert lib.whisper_tokenize(ctx.ptr, b"hello world", buf, 4) == -11
    a.free(); b.free()


def test_language_auto_detect_equals_oracle(prod_lib, ref_lib):
    """whisper_lang_auto_detect_with_state (wc:4037-4110): encode, decode [sot], softmax over the language tokens"""
    pcm = synth.synth_audio(11.0, seed=3)
    res = []
    for lib, gpu in ((ref_lib, False), (prod_lib, True)):
        ctx = _ctx(lib, gpu)
        st = ctx.create_state()
        st.pcm_to_mel(pcm, threads=4)
        res.append(st.lang_detect(0, 4))
        with pytest.raises(api.WhisperError):
            st.lang_detect(60_000, 4)   # offset past the end of the audio -> negative return
        st.free(); ctx.free()
    (id_r, p_r), (id_p, p_p) = res
    assert abs(float(p_r.sum()) - 1.0) < 1e-4 and abs(float(p_p.sum()) - 1.0) < 1e-4
    np.testing.assert_allclose(p_p, p_r, atol=2e-2, rtol=5e-2)
    top = np.sort(p_r)[-2:]
    if top[1] - top[0] > 0.02:
        assert id_p == id_r


def test_set_mel_then_full_equals_pcm_path(prod_lib):
    """whisper_set_mel_with_state (wc:3891-3913): a mel computed by pcm_to_mel, fed back through set_mel, encodes identically;
    a wrong band count is refused with -1"""
    ctx = _ctx(prod_lib, True)
    st = ctx.create_state()
    st.pcm_to_mel(synth.synth_audio(11.0, seed=3), threads=4)
    mel, _ = prod_mel(prod_lib, st)
    st.encode(0, 4)
    st.decode([ctx.token_sot()], 0, 4)
    want = st.get_logits()
    st2 = ctx.create_state()
    st2.set_mel(mel)
    st2.encode(0, 4)
    st2.decode([ctx.token_sot()], 0, 4)
    np.testing.assert_array_equal(st2.get_logits(), want)
    with pytest.raises(api.WhisperError):
        st2.set_mel(mel[:-1])
    st.free(); st2.free(); ctx.free()


def test_full_parallel_equals_oracle(prod_lib, ref_lib):
    """whisper_full_parallel (wc:7736-7864): two equal sample spans, one state each, times of the second span offset and
    clamped; here the spans advance as one lock-step batch on the GPU"""
    pcm = synth.synth_audio(65.0, seed=28)
    out = []
    for lib in (ref_lib, prod_lib):
        for name, res, args in (("whisper_init_from_buffer_with_params", C.c_void_p, [C.c_void_p, C.c_size_t, api.whisper_context_params]),
                                ("whisper_full_parallel", C.c_int, [C.c_void_p, api.whisper_full_params, C.POINTER(C.c_float), C.c_int, C.c_int]),
                                ("whisper_full_n_segments", C.c_int, [C.c_void_p]),
                                ("whisper_full_get_segment_t0", C.c_int64, [C.c_void_p, C.c_int]),
                                ("whisper_full_get_segment_t1", C.c_int64, [C.c_void_p, C.c_int]),
                                ("whisper_full_get_segment_text", C.c_char_p, [C.c_void_p, C.c_int]),
                                ("whisper_full_n_tokens", C.c_int, [C.c_void_p, C.c_int]),
                                ("whisper_full_get_token_id", C.c_int32, [C.c_void_p, C.c_int, C.c_int])):
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        blob = model_blob("mini", RICH)
        cp, _keep = WhisperContextParameters(use_gpu=lib is prod_lib).to_c(lib)
        ctx = lib.whisper_init_from_buffer_with_params(blob, len(blob), cp)
        assert ctx
        p = FullParams(lib, SamplingStrategy.GREEDY, best_of=1)
        p.set_n_threads(4).set_language("en").set_temperature_inc(0.0)
        assert lib.whisper_full_parallel(ctx, p.fp, pcm.ctypes.data_as(C.POINTER(C.c_float)), len(pcm), 2) == 0
        segs = []
        for i in range(lib.whisper_full_n_segments(ctx)):
            segs.append((lib.whisper_full_get_segment_t0(ctx, i), lib.whisper_full_get_segment_t1(ctx, i), lib.whisper_full_get_segment_text(ctx, i),
                         [lib.whisper_full_get_token_id(ctx, i, j) for j in range(lib.whisper_full_n_tokens(ctx, i))]))
        out.append(segs)
        lib.whisper_free(ctx)
    ref, got = out
    # the second span of this clip runs into a near-tie late in its second window (the reference's own ISA builds disagree on such
    # cases): the engines must agree on a long prefix, including the offset / clamped times of the second span ...
    same = next((k for k, (x, y) in enumerate(zip(ref, got)) if x != y), min(len(ref), len(got)))
    assert len(ref) > 1 and same >= 15 and any(s[0] >= 3250 for s in got[:same])
    # ... and whisper_full_parallel must be exactly the per-span transcriptions merged by the reference's rule (wc:7809-7822)
    ctx = _ctx(prod_lib, True)
    p = FullParams(prod_lib, SamplingStrategy.GREEDY, best_of=1)
    p.set_n_threads(4).set_language("en").set_temperature_inc(0.0)
    half, want = len(pcm) // 2, []
    for i in range(2):
        st = ctx.create_state()
        assert st.full(p, pcm[i * half:(i + 1) * half]) == 0
        off = 100 * (i * half) // 16000
        for sg in st.segments():
            t0, t1 = sg.t0 + off, sg.t1 + off
            if i > 0 and want:
                t0 = max(t0, want[-1][1])
            want.append((t0, t1, sg.text, [t["id"] for t in sg.tokens]))
        st.free()
    ctx.free()
    assert got == want
