"""SURVEY 8(f).4 — the VAD front end (whisper.cpp:4361-5480, 6615-6793, 7882-7960).

CPU (`not gpu`): the model loader and the speech-segment rules (whisper_vad_segments_from_probs) are host logic; they are checked against
the reference on the oracle's own probabilities and on random / adversarial sequences over a grid of parameters — exact equality.
GPU: the three kernels' probabilities against the oracle, and whisper_full / whisper_full_parallel with params.vad end to end (filtered
samples, transcript, and segment times mapped back through the table)."""
import ctypes as C

import numpy as np
import pytest

import vad_helpers as vh
from conftest import model_blob
from whisper_rust_b200 import FullParams, SamplingStrategy, synth

RICH = synth.InitRecipe(te_std=0.2, pe_std=0.3, gain=3.0, ts_scale=1.6)


def _set_probs(prod_lib, vad, p):
    prod_lib.whisper_b200_test_vad_set_probs.restype = C.c_int
    prod_lib.whisper_b200_test_vad_set_probs.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    a = np.ascontiguousarray(p, dtype=np.float32)
    assert prod_lib.whisper_b200_test_vad_set_probs(vad.ptr, a.ctypes.data, len(a)) == len(a)


def _params(lib, **kw):
    p = vh.declare(lib).whisper_vad_default_params()
    for k, v in kw.items():
        setattr(p, k, v)
    return p


PARAM_GRID = [dict(), dict(threshold=0.3), dict(threshold=0.7, min_silence_duration_ms=300), dict(min_speech_duration_ms=1000),
              dict(max_speech_duration_s=2.0), dict(max_speech_duration_s=1.0, min_silence_duration_ms=50), dict(speech_pad_ms=400),
              dict(speech_pad_ms=0, min_speech_duration_ms=0, min_silence_duration_ms=0), dict(threshold=0.1, max_speech_duration_s=3.5)]


def test_segment_rules_equal_the_reference(prod_lib, ref_lib):
    """whisper_vad_segments_from_probs is integer / float host logic: same probabilities in -> identical segments out"""
    ref = vh.Vad(ref_lib)
    prod = vh.Vad(prod_lib)          # loads on the host; no device is touched until detect_speech
    rng = np.random.default_rng(5)
    seqs = [ref.probs(synth.gated_audio(30.0, seed=5)), ref.probs(synth.gated_audio(65.0, seed=9))]
    for n in (0, 1, 7, 300, 2000):
        seqs.append(rng.uniform(0, 1, n).astype(np.float32))
    sm = np.clip(np.convolve(rng.uniform(0, 1, 3000), np.ones(25) / 25, mode="same") * 1.6 - 0.3, 0, 1).astype(np.float32)   # long runs
    seqs += [sm, (sm > 0.5).astype(np.float32), np.ones(500, dtype=np.float32), np.zeros(500, dtype=np.float32),
             np.tile(np.array([0.9] * 40 + [0.4] * 4, dtype=np.float32), 30)]
    n_seg = 0
    for p in seqs:
        _ref_set_probs(ref_lib, ref, p)
        _set_probs(prod_lib, prod, p)
        for kw in PARAM_GRID:
            want = ref.segments(_params(ref_lib, **kw))
            got = prod.segments(_params(prod_lib, **kw))
            assert got == want, (len(p), kw)
            n_seg += len(want)
    assert n_seg > 200
    ref.free(); prod.free()


def test_segment_rules_equal_the_reference_on_random_runs_and_parameters(prod_lib, ref_lib):
    """the same, fuzzed: telegraph-like probability tracks (speech / silence runs of random lengths, levels hovering around the two
    thresholds, single-window spikes and drop-outs) under randomly drawn parameters, incl. the max-speech splitting rules"""
    ref, prod = vh.Vad(ref_lib), vh.Vad(prod_lib)
    rng = np.random.default_rng(17)
    n_seg, n_cases, n_split = 0, 0, 0
    for k in range(120):
        n = int(rng.integers(1, 700))
        track = np.empty(0, dtype=np.float32)
        while len(track) < n:
            speech = rng.random() < 0.55
            run = int(rng.integers(1, 4)) if rng.random() < 0.2 else int(rng.integers(4, 120))
            level = rng.uniform(0.45, 1.0) if speech else rng.uniform(0.0, 0.55)
            seg = level + rng.normal(0, rng.choice([0.0, 0.02, 0.1]), run)
            track = np.concatenate([track, np.clip(seg, 0, 1).astype(np.float32)])
        p = track[:n]
        _ref_set_probs(ref_lib, ref, p)
        _set_probs(prod_lib, prod, p)
        for _ in range(4):
            kw = dict(threshold=float(rng.choice([0.2, 0.35, 0.5, 0.5, 0.65, 0.9])),
                      min_speech_duration_ms=int(rng.choice([0, 32, 100, 250, 600])),
                      min_silence_duration_ms=int(rng.choice([0, 32, 100, 300, 1000])),
                      max_speech_duration_s=float(rng.choice([0.5, 1.0, 2.5, 6.0, 1e9])),
                      speech_pad_ms=int(rng.choice([0, 30, 100, 500])),
                      samples_overlap=float(rng.choice([0.0, 0.1, 0.5])))
            want = ref.segments(_params(ref_lib, **kw))
            got = prod.segments(_params(prod_lib, **kw))
            assert got == want, (k, n, kw)
            n_seg += len(want)
            n_cases += 1
            n_split += int(kw["max_speech_duration_s"] < 3.0 and len(want) > 1)
    print("vad fuzz:", n_cases, "cases", n_seg, "segments", n_split, "split cases")
    assert n_cases == 480 and n_seg > 500 and n_split > 20
    ref.free(); prod.free()


def _ref_set_probs(ref_lib, ref, p):
    """give the reference an arbitrary probability sequence: run detect_speech on len(p) windows of silence, then overwrite the
    probabilities it holds (whisper_vad_probs returns its own buffer)"""
    n = len(p)
    z = np.zeros(max(512 * n, 1), dtype=np.float32)
    assert ref_lib.whisper_vad_detect_speech(ref.ptr, z.ctypes.data, 512 * n)
    assert ref_lib.whisper_vad_n_probs(ref.ptr) == n
    if n:
        np.ctypeslib.as_array(ref_lib.whisper_vad_probs(ref.ptr), shape=(n,))[:] = p


def test_vad_loader_rejects_what_the_reference_rejects(prod_lib, ref_lib, tmp_path):
    vh.declare(prod_lib); vh.declare(ref_lib)
    good = synth.vad_model_bytes()
    cases = {"bad_magic": b"\x00" * 4 + good[4:], "truncated": good[:len(good) // 2], "empty": b""}
    for name, blob in cases.items():
        f = tmp_path / f"{name}.bin"
        f.write_bytes(blob)
        for lib in (prod_lib, ref_lib):
            if name == "truncated" and lib is ref_lib:
                continue   # the reference reads past a truncated tensor without noticing; this loader reports it
            assert not lib.whisper_vad_init_from_file_with_params(str(f).encode(), lib.whisper_vad_default_context_params()), name
    assert not prod_lib.whisper_vad_init_from_file_with_params(b"/nonexistent/vad.bin", prod_lib.whisper_vad_default_context_params())


@pytest.mark.gpu
@pytest.mark.parametrize("seconds,seed", [(30.0, 5), (65.0, 9), (11.0, 3), (0.02, 1), (47.3, 12)])
def test_vad_probabilities_match_oracle(prod_lib, ref_lib, seconds, seed):
    pcm = synth.gated_audio(seconds, seed=seed)
    ref, prod = vh.Vad(ref_lib), vh.Vad(prod_lib)
    a, b = ref.probs(pcm), prod.probs(pcm)
    assert a.shape == b.shape and len(a) == -(-len(pcm) // 512)
    err = float(np.max(np.abs(a - b))) if len(a) else 0.0
    print(f"VAD probabilities over {len(a)} windows: max abs err {err:.2e}; range {a.min() if len(a) else 0:.2f}..{a.max() if len(a) else 0:.2f}")
    assert err < 2e-3            # f16 conv inputs / f32 accumulation on both sides, another summation order; sigmoid output in [0, 1]
    # segments: identical wherever no probability sits within that error of a threshold
    safe = not np.any((np.abs(a - 0.5) < 2 * err + 1e-6) | (np.abs(a - 0.35) < 2 * err + 1e-6))
    if safe:
        assert ref.segments() == prod.segments()
    ref.free(); prod.free()


def _full_vad(lib, gpu, pcm, n_processors=1, vad=True):
    ctx, keep = vh.default_state_context(lib, model_blob("mini", RICH), gpu)
    p = FullParams(lib, SamplingStrategy.GREEDY, best_of=1)
    p.set_n_threads(8).set_language("en").set_temperature_inc(0.0)
    p.fp.vad = vad
    path = vh.vad_model_path().encode()
    p.fp.vad_model_path = path
    a = np.ascontiguousarray(pcm, dtype=np.float32)
    fn = lib.whisper_full_parallel if n_processors > 1 else lib.whisper_full
    args = (ctx, p.fp, a.ctypes.data_as(C.POINTER(C.c_float)), len(a)) + ((n_processors,) if n_processors > 1 else ())
    r = fn(*args)
    segs = vh.default_state_segments(lib, ctx)
    lib.whisper_free(ctx)
    return r, segs


# clips on which the reference's builds and the CUDA path agree end to end ((65.0, 9), (65.0, 11), (65.0, 21), (65.0, 23) flip a near-tie
# token in the transcription of the filtered audio between the reference's OWN two builds, (65.0, 14) between the reference and B200: the
# VAD part - probabilities, segments, filtered samples - is identical on all of them, see test_vad_probabilities_match_oracle)
@pytest.mark.gpu
@pytest.mark.parametrize("seconds,seed", [(30.0, 5), (47.3, 12), (47.3, 7)])
def test_whisper_full_with_vad_matches_oracle(prod_lib, ref_lib, seconds, seed):
    """params.vad through whisper_full (default state): speech segments filtered, transcribed, and the segment times mapped back"""
    pcm = synth.gated_audio(seconds, seed=seed)
    ra, sa = _full_vad(ref_lib, False, pcm)
    rb, sb = _full_vad(prod_lib, True, pcm)
    assert ra == rb == 0
    assert sb and [(t0, t1) for t0, t1, _, _ in sa] == [(t0, t1) for t0, t1, _, _ in sb]
    assert [ids for *_, ids in sa] == [ids for *_, ids in sb]
    # and the mapping really moved the times: without VAD the same call gives other segment times
    _, plain = _full_vad(prod_lib, True, pcm, vad=False)
    assert [(t0, t1) for t0, t1, _, _ in plain] != [(t0, t1) for t0, t1, _, _ in sb]


@pytest.mark.gpu
def test_whisper_full_with_vad_on_silence_and_with_state_ignores_vad(prod_lib, ref_lib):
    quiet = (synth.synth_audio(11.0, seed=3) * 1e-4).astype(np.float32)
    ra, sa = _full_vad(ref_lib, False, quiet)
    rb, sb = _full_vad(prod_lib, True, quiet)
    assert (ra, sa) == (rb, sb) == (0, [])                       # no speech: empty result, return 0 (wc:7726-7729)
    # whisper_full_with_state never reads params.vad (wc:6795-7711): same transcript with the flag on or off
    from whisper_rust_b200 import WhisperContext, WhisperContextParameters
    ctx = WhisperContext.new_from_buffer_with_params(model_blob("mini", RICH), WhisperContextParameters(use_gpu=True), lib=prod_lib)
    st = ctx.create_state()
    pcm = synth.gated_audio(30.0, seed=5)
    out = []
    for flag in (False, True):
        p = FullParams(prod_lib, SamplingStrategy.GREEDY, best_of=1)
        p.set_n_threads(8).set_language("en").set_temperature_inc(0.0)
        p.fp.vad = flag
        assert st.full(p, pcm) == 0
        out.append([(s.t0, s.t1, [t["id"] for t in s.tokens]) for s in st.segments()])
    assert out[0] == out[1] and out[0]
    st.free(); ctx.free()


@pytest.mark.gpu
def test_whisper_full_parallel_with_vad_matches_oracle(prod_lib, ref_lib):
    pcm = synth.gated_audio(65.0, seed=9)
    ra, sa = _full_vad(ref_lib, False, pcm, n_processors=2)
    rb, sb = _full_vad(prod_lib, True, pcm, n_processors=2)
    assert ra == rb == 0 and sb
    assert [(t0, t1, ids) for t0, t1, _, ids in sa] == [(t0, t1, ids) for t0, t1, _, ids in sb]
