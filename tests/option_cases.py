"""The FullParams option surface of whisper_full_with_state (wc:6795-7711, rs/whisper_params.rs) as a table of cases that both the
oracle and the product run through the identical call sequence.  tools/option_stability.py runs the table on the reference's two ISA
builds and prints which cases they agree on; only those are listed in STABLE (a near-tie argmax is decided by summation order, which
differs even between two builds of the reference itself: DESIGN.md section 2)."""
from whisper_rust_b200 import FullParams, SamplingStrategy, synth

RICH = synth.InitRecipe(te_std=0.2, pe_std=0.3, gain=3.0, ts_scale=1.6)


def _base(lib, strategy=SamplingStrategy.GREEDY, **kw):
    p = FullParams(lib, strategy, **kw)
    p.set_n_threads(8).set_language("en").set_temperature_inc(0.0)
    return p


BASE = synth.InitRecipe(te_std=0.2)


def model_of(name):
    """(arch, recipe) of a case: mini/RICH unless the case names another"""
    spec = CASES[name]
    return spec.get("arch", "mini"), spec.get("recipe", RICH)


def run_case(name, lib, ctx):
    """returns a list of per-call results [(ret, segments, lang_id)]; every call uses a fresh FullParams like a Rust caller would"""
    spec = CASES[name]
    st = ctx.create_state()
    out = []
    for call in spec["calls"]:
        p = _base(lib, best_of=1)
        call["opts"](p, ctx)
        pcm = synth.synth_audio(call["seconds"], seed=call["seed"])
        try:
            r = st.full(p, pcm)
        except Exception as e:      # WhisperError carries the code
            r = getattr(e, "code", -999)
        out.append((r, st.segments(), st.full_lang_id()))
    st.free()
    return out


def _c(seconds, seed, opts):
    return dict(seconds=seconds, seed=seed, opts=opts)


def _prompt_tokens(p, ctx):
    p.set_tokens(ctx.tokenize(" t300t301 t303", 16))


CASES = {
    # prompt assembly (wc:6907-6965, 7073-7085)
    "initial_prompt": dict(calls=[_c(30.0, 26, lambda p, c: p.set_initial_prompt(" t300t301 t303t400"))]),
    "prompt_tokens": dict(calls=[_c(30.0, 26, _prompt_tokens)]),
    "carry_over_no_context_false": dict(calls=[_c(30.0, 26, lambda p, c: p.set_no_context(False)),
                                               _c(47.3, 26, lambda p, c: p.set_no_context(False)),
                                               _c(11.0, 3, lambda p, c: p.set_no_context(False))]),
    "carry_over_then_reset": dict(calls=[_c(47.3, 26, lambda p, c: p.set_no_context(False)),
                                         _c(30.0, 26, lambda p, c: p.set_no_context(True))]),
    # decoding limits and segment shaping
    "no_timestamps": dict(calls=[_c(47.3, 26, lambda p, c: p.set_no_timestamps(True))]),
    "single_segment": dict(calls=[_c(47.3, 26, lambda p, c: p.set_single_segment(True))]),
    "max_tokens_12": dict(calls=[_c(47.3, 26, lambda p, c: p.set_max_tokens(12))]),
    "offset_duration": dict(calls=[_c(65.0, 28, lambda p, c: p.set_offset_ms(7000).set_duration_ms(31000))]),
    "offset_only": dict(calls=[_c(47.3, 26, lambda p, c: p.set_offset_ms(12340))]),
    "audio_ctx_768": dict(calls=[_c(11.0, 3, lambda p, c: p.set_audio_ctx(768))]),
    "max_initial_ts_0": dict(calls=[_c(30.0, 26, lambda p, c: p.set_max_initial_ts(0.0))]),
    "no_suppress_blank": dict(calls=[_c(30.0, 26, lambda p, c: p.set_suppress_blank(False))]),
    # suppression rules (wc:6197-6260)
    "suppress_nst": dict(calls=[_c(47.3, 26, lambda p, c: p.set_suppress_nst(True))]),
    "suppress_regex": dict(calls=[_c(47.3, 26, lambda p, c: p.set_suppress_regex(r" ?t[0-9]*[02468]"))]),
    "tdrz_enable": dict(calls=[_c(30.0, 26, lambda p, c: p.set_tdrz_enable(True))]),
    # task / language (wc:6941-6952, 6815-6840)
    "translate": dict(calls=[_c(30.0, 26, lambda p, c: p.set_translate(True))]),
    "language_de": dict(calls=[_c(30.0, 26, lambda p, c: p.set_language("de"))]),
    "language_auto": dict(calls=[_c(30.0, 26, lambda p, c: p.set_language("auto"))]),
    "detect_language_only": dict(calls=[_c(30.0, 26, lambda p, c: p.set_language("auto").set_detect_language(True))]),
    # thresholds that gate the fallback ladder without sampling (temperature_inc = 0 keeps a single pass; wc:7484-7544)
    "no_speech_thold_1": dict(calls=[_c(30.0, 26, lambda p, c: p.set_no_speech_thold(1.0).set_logprob_thold(0.0))]),
    "length_penalty": dict(calls=[_c(30.0, 26, lambda p, c: p.set_length_penalty(0.7))]),
    # callbacks that steer control flow (wc:6987-7004, 2471)
    "encoder_begin_false": dict(calls=[_c(30.0, 26, lambda p, c: p.set_encoder_begin_callback(lambda a, b, u: False))]),
    "abort_always": dict(calls=[_c(30.0, 26, lambda p, c: p.set_abort_callback(lambda u: True))]),
    # alternates on other clips (kept when the first choice is a near-tie case on the reference's own builds)
    "carry_over_b": dict(calls=[_c(11.0, 3, lambda p, c: p.set_no_context(False)), _c(30.0, 26, lambda p, c: p.set_no_context(False))]),
    "carry_over_c": dict(calls=[_c(30.0, 26, lambda p, c: p.set_no_context(False)), _c(30.0, 26, lambda p, c: p.set_no_context(False))]),
    "carry_over_d": dict(calls=[_c(65.0, 28, lambda p, c: p.set_no_context(False)), _c(11.0, 3, lambda p, c: p.set_no_context(False))]),
    "no_timestamps_b": dict(calls=[_c(30.0, 26, lambda p, c: p.set_no_timestamps(True))]),
    "no_timestamps_c": dict(calls=[_c(11.0, 3, lambda p, c: p.set_no_timestamps(True))]),
    "no_timestamps_d": dict(calls=[_c(65.0, 28, lambda p, c: p.set_no_timestamps(True))]),
    "offset_duration_b": dict(calls=[_c(47.3, 26, lambda p, c: p.set_offset_ms(3000).set_duration_ms(20000))]),
    "offset_duration_c": dict(calls=[_c(65.0, 28, lambda p, c: p.set_offset_ms(30000).set_duration_ms(12000))]),
    "offset_duration_d": dict(calls=[_c(47.3, 25, lambda p, c: p.set_offset_ms(1500).set_duration_ms(33000))]),
    "offset_only_b": dict(calls=[_c(65.0, 28, lambda p, c: p.set_offset_ms(20000))]),
    "offset_only_c": dict(calls=[_c(47.3, 25, lambda p, c: p.set_offset_ms(8000))]),
    "offset_only_d": dict(calls=[_c(30.0, 26, lambda p, c: p.set_offset_ms(4000))]),
    "audio_ctx_b": dict(calls=[_c(30.0, 26, lambda p, c: p.set_audio_ctx(1024))]),
    "audio_ctx_c": dict(calls=[_c(11.0, 3, lambda p, c: p.set_audio_ctx(600))]),
    "audio_ctx_d": dict(calls=[_c(5.0, 5, lambda p, c: p.set_audio_ctx(256))]),
    "suppress_regex_b": dict(calls=[_c(30.0, 26, lambda p, c: p.set_suppress_regex(r" ?t[0-9]*[02468]"))]),
    "suppress_regex_c": dict(calls=[_c(47.3, 26, lambda p, c: p.set_suppress_regex(r"t[0-9]*7"))]),
    "suppress_regex_d": dict(calls=[_c(11.0, 3, lambda p, c: p.set_suppress_regex(r" ?t[0-9]*[13579]"))]),
    "no_timestamps_tiny": dict(arch="tiny", recipe=BASE, calls=[_c(30.0, 3, lambda p, c: p.set_no_timestamps(True))]),
    "no_timestamps_tiny_b": dict(arch="tiny", recipe=BASE, calls=[_c(47.3, 4, lambda p, c: p.set_no_timestamps(True))]),
    # heuristic token timestamps ride on most of the above (wc:8390-8615)
    "token_timestamps_offset": dict(calls=[_c(47.3, 26, lambda p, c: p.set_token_timestamps(True).set_offset_ms(5000).set_max_len(24))]),
}


# cases on which the reference's AVX2 and -march=native builds agree (tools/option_stability.py, run in the build container);
# the others (long free-running outputs of the RICH recipe: carry_over_no_context_false, no_timestamps on mini, offset_only,
# offset_duration on 65 s, suppress_regex on 47.3 s ...) flip a near-tie argmax between the reference's own two builds
STABLE = ['initial_prompt', 'prompt_tokens', 'carry_over_then_reset', 'carry_over_c', 'carry_over_d', 'single_segment', 'max_tokens_12',
          'max_initial_ts_0', 'no_suppress_blank', 'suppress_nst', 'suppress_regex_b', 'suppress_regex_c', 'suppress_regex_d', 'tdrz_enable',
          'translate', 'language_de', 'language_auto', 'detect_language_only', 'no_speech_thold_1', 'length_penalty', 'encoder_begin_false',
          'abort_always', 'token_timestamps_offset', 'offset_duration_b', 'offset_duration_c', 'offset_only_d', 'audio_ctx_b', 'audio_ctx_c',
          'audio_ctx_d', 'no_timestamps_tiny', 'no_timestamps_tiny_b']
