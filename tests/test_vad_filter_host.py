"""The host half of the VAD pre-pass without a device: speech segments -> the filtered sample buffer whisper_full transcribes and the
processed-time -> original-time table (csrc/vad_api.cu: vad_filter_segments, vad_map_time_table) against the reference's own static
whisper_vad / map_processed_to_original_time (whisper.cpp:6615-6793, 7882-7925; reached through oracle/ref_probe.cpp).  The reference
runs its Silero graph on the CPU; the product is handed the reference's probabilities (the device kernels that produce them are checked
on the GPU in tests/test_vad.py), so everything compared here is integer / copy logic: exact equality."""
import ctypes as C

import numpy as np
import pytest

import vad_helpers as vh
from whisper_rust_b200 import FullParams, SamplingStrategy, api, synth

SIG_TAIL = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.c_void_p, C.c_void_p, C.c_int]

# (seconds, seed, gating period, duty); the last two reach the end of the clip in speech (the reference sizes the buffer with
# min(end, n - 1) and copies min(end, n): the product keeps the size and copies what fits)
CLIPS = [(30.0, 5, 4.0, 0.55), (47.3, 12, 4.0, 0.55), (65.0, 9, 7.0, 0.4), (11.0, 3, 2.0, 0.5), (20.0, 7, 12.0, 0.8), (8.0, 2, 30.0, 1.0),
         (16.0, 4, 6.0, 0.9)]
PARAMS = [dict(), dict(samples_overlap=0.0), dict(samples_overlap=0.5), dict(speech_pad_ms=0), dict(speech_pad_ms=400, samples_overlap=0.3),
          dict(threshold=0.3, min_silence_duration_ms=300), dict(max_speech_duration_s=1.5), dict(max_speech_duration_s=0.7, samples_overlap=0.2),
          dict(min_speech_duration_ms=1500), dict(threshold=0.99)]


@pytest.fixture(scope="module")
def sides(prod_lib, ref_lib):
    vh.declare(prod_lib); vh.declare(ref_lib)
    ref_lib.ref_probe_vad_filter.restype = C.c_int
    ref_lib.ref_probe_vad_filter.argtypes = [C.c_void_p, api.whisper_full_params] + SIG_TAIL
    prod_lib.whisper_b200_test_vad_filter.restype = C.c_int
    prod_lib.whisper_b200_test_vad_filter.argtypes = [C.c_void_p, api.whisper_vad_params] + SIG_TAIL
    prod_lib.whisper_b200_test_vad_set_probs.restype = C.c_int
    prod_lib.whisper_b200_test_vad_set_probs.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    blob = synth.model_bytes("micro", synth.InitRecipe(), with_tensors=False)      # whisper_vad only needs a context with a default state
    ctx, keep = vh.default_state_context(ref_lib, blob, False)
    ref_vad, prod_vad = vh.Vad(ref_lib), vh.Vad(prod_lib)                            # the product's context loads on the host
    path = vh.vad_model_path().encode()
    grid = np.arange(-40, 9000, 7, dtype=np.int64)

    probs_of = {}

    def run(pcm, kw):
        n, cap, tcap = len(pcm), len(pcm) + 64, 8192
        if pcm.ctypes.data not in probs_of:                                          # the reference's probabilities, once per clip
            probs_of.clear()
            probs_of[pcm.ctypes.data] = ref_vad.probs(pcm)
        p = FullParams(ref_lib, SamplingStrategy.GREEDY, best_of=1)
        p.fp.vad, p.fp.vad_model_path = True, path
        vp = ref_lib.whisper_vad_default_params()
        for k, v in kw.items():
            setattr(vp, k, v)
        p.fp.vad_params = vp
        out = []
        for side in ("ref", "prod"):
            buf = np.full(cap, np.nan, dtype=np.float32)
            tab = np.zeros((tcap, 2), dtype=np.int64)
            nt = C.c_int(-1)
            mo = np.zeros_like(grid)
            if side == "ref":
                r = ref_lib.ref_probe_vad_filter(ctx, p.fp, pcm.ctypes.data, n, buf.ctypes.data, cap, tab.ctypes.data, tcap, C.byref(nt),
                                                 grid.ctypes.data, mo.ctypes.data, len(grid))
            else:
                probs = probs_of[pcm.ctypes.data]
                assert prod_lib.whisper_b200_test_vad_set_probs(prod_vad.ptr, probs.ctypes.data, len(probs)) == len(probs)
                r = prod_lib.whisper_b200_test_vad_filter(prod_vad.ptr, vp, pcm.ctypes.data, n, buf.ctypes.data, cap, tab.ctypes.data, tcap,
                                                          C.byref(nt), grid.ctypes.data, mo.ctypes.data, len(grid))
            assert r >= 0 and 0 <= nt.value <= tcap
            out.append((r, buf[:r].copy(), tab[:nt.value].copy(), mo))
        return out

    yield run
    ref_vad.free(); prod_vad.free()
    ref_lib.whisper_free(ctx)


@pytest.mark.parametrize("seconds,seed,period,duty", CLIPS)
def test_filtered_samples_and_time_table_equal_the_reference(sides, seconds, seed, period, duty):
    pcm = np.ascontiguousarray(synth.gated_audio(seconds, seed=seed, period=period, duty=duty))
    n_speech, n_points = 0, 0
    for kw in (PARAMS if seconds <= 20.0 else PARAMS[::2]):                        # each case re-runs the reference's Silero graph on the CPU
        (ra, fa, ta, ma), (rb, fb, tb, mb) = sides(pcm, kw)
        assert ra == rb, (kw, ra, rb)
        # the reference's copy of the final segment can run one float past its buffer; inside the buffer the two must agree bit for bit
        np.testing.assert_array_equal(fa.view(np.uint32), fb.view(np.uint32), err_msg=str(kw))
        np.testing.assert_array_equal(ta, tb, err_msg=str(kw))
        np.testing.assert_array_equal(ma, mb, err_msg=str(kw))
        n_speech += int(ra > 0)
        n_points += len(ta)
    assert n_speech >= 3 and n_points >= 40
