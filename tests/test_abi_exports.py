"""The drop-in boundary without a GPU: libwhisper_b200.so loads, exports every symbol include/*.h declares, keeps the
reference's struct layouts and defaults, and fails loudly (NULL context, logged reason) where it would need a device.
Known answers come from tests/golden/known_answers.json, dumped from the reference itself (tests/golden/make_golden.py)."""
import ctypes as C
import json
import re
from pathlib import Path

import numpy as np
import pytest

from conftest import GOLDEN, ROOT
from whisper_rust_b200 import api, synth, WhisperContext, WhisperContextParameters

KA = json.loads((GOLDEN / "known_answers.json").read_text())


def declared_symbols():
    names = []
    for h in sorted((ROOT / "include").glob("*.h")):
        txt = re.sub(r"/\*.*?\*/", "", h.read_text(), flags=re.S)
        txt = re.sub(r"//[^\n]*", "", txt)
        txt = re.sub(r"#\s*define[^\n]*", "", txt)
        for m in re.finditer(r"WHISPER_API\s+[^;{]*?\b(\w+)\s*\(", txt):
            names.append((h.name, m.group(1)))
    return names


def test_headers_declare_the_reference_abi():
    names = {n for _, n in declared_symbols()}
    # the must-export set the Rust crate binds
    for n in ("whisper_init_from_file_with_params_no_state", "whisper_init_from_buffer_with_params_no_state", "whisper_init_state",
              "whisper_free", "whisper_free_state", "whisper_full_default_params", "whisper_context_default_params",
              "whisper_full_with_state", "whisper_pcm_to_mel_with_state", "whisper_set_mel_with_state", "whisper_encode_with_state",
              "whisper_decode_with_state", "whisper_lang_auto_detect_with_state", "whisper_get_logits_from_state",
              "whisper_full_n_segments_from_state", "whisper_full_lang_id_from_state", "whisper_full_get_segment_t0_from_state",
              "whisper_full_get_segment_t1_from_state", "whisper_full_get_segment_text_from_state",
              "whisper_full_get_segment_speaker_turn_next_from_state", "whisper_full_n_tokens_from_state",
              "whisper_full_get_token_text_from_state", "whisper_full_get_token_id_from_state", "whisper_full_get_token_data_from_state",
              "whisper_full_get_token_p_from_state", "whisper_n_len_from_state", "whisper_n_vocab", "whisper_n_text_ctx",
              "whisper_n_audio_ctx", "whisper_is_multilingual", "whisper_model_n_vocab", "whisper_model_n_mels", "whisper_model_ftype",
              "whisper_model_type", "whisper_model_type_readable", "whisper_tokenize", "whisper_token_to_str", "whisper_token_eot",
              "whisper_token_sot", "whisper_token_beg", "whisper_token_lang", "whisper_lang_id", "whisper_lang_max_id", "whisper_lang_str",
              "whisper_lang_str_full", "whisper_log_set", "whisper_print_system_info", "whisper_print_timings", "whisper_reset_timings",
              "whisper_full_parallel", "ggml_log_set", "ggml_cpu_has_avx", "ggml_cpu_has_avx2", "ggml_cpu_has_fma", "ggml_cpu_has_f16c"):
        assert n in names, f"{n} is not declared in include/*.h"


def test_library_exports_every_declared_symbol():
    lib = api.load_library()
    missing = [f"{h}:{n}" for h, n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing


def test_struct_layouts_match_the_reference_header():
    # measured from the vendored header with gcc
    assert C.sizeof(api.whisper_full_params) == 296
    assert C.sizeof(api.whisper_context_params) == 48
    assert C.sizeof(api.whisper_token_data) == 56
    fp = api.whisper_full_params
    assert (fp.strategy.offset, fp.n_threads.offset, fp.language.offset, fp.greedy.offset, fp.beam_search.offset) == (0, 4, 96, 136, 140)
    assert (fp.new_segment_callback.offset, fp.grammar_rules.offset, fp.grammar_penalty.offset, fp.vad.offset) == (152, 232, 256, 260)
    assert (fp.vad_model_path.offset, fp.vad_params.offset) == (264, 272)
    cp = api.whisper_context_params
    assert (cp.dtw_token_timestamps.offset, cp.dtw_aheads_preset.offset, cp.dtw_n_top.offset, cp.dtw_aheads.offset, cp.dtw_mem_size.offset) == (8, 12, 16, 24, 40)
    td = api.whisper_token_data
    assert (td.t0.offset, td.t_dtw.offset, td.vlen.offset) == (24, 40, 48)


@pytest.mark.parametrize("name,strategy", [("greedy", 0), ("beam", 1)])
def test_default_params_equal_the_reference(name, strategy):
    lib = api.load_library()
    fp = lib.whisper_full_default_params(strategy)
    want = KA["defaults"][name]
    for k, v in want.items():
        if "." in k:
            a, b = k.split(".")
            got = getattr(getattr(fp, a), b)
        else:
            got = getattr(fp, k)
            if isinstance(got, bytes):
                got = got.decode()
        if k == "n_threads":   # min(4, hardware_concurrency) (wc:5918)
            assert 1 <= got <= 4
            continue
        assert got == pytest.approx(v), k
    cp = lib.whisper_context_default_params()
    for k, v in KA["defaults"]["context"].items():
        assert getattr(cp, k) == v, k


def test_language_table_equals_the_reference():
    lib = api.load_library()
    assert lib.whisper_lang_max_id() + 1 == len(KA["langs"]) == 100
    for i, (code, name) in enumerate(KA["langs"]):
        assert lib.whisper_lang_str(i).decode() == code
        assert lib.whisper_lang_str_full(i).decode() == name
        assert lib.whisper_lang_id(code.encode()) == i
        assert lib.whisper_lang_id(name.encode()) == i
    assert lib.whisper_lang_id(b"klingon") == -1           # wc:4009-4010
    assert lib.whisper_print_system_info() is not None


def test_no_device_means_loud_failure_not_a_cpu_path():
    """Without a CUDA device context creation returns NULL and logs why; garbage / missing model files are rejected
    like the reference's own error-path tests (src/whisper_ctx_wrapper.rs:490-513,598-614)."""
    import torch
    lib = api.load_library()
    logs = []
    cb = api.LOG_CB(lambda lvl, msg, ud: logs.append((lvl, msg)))
    lib.whisper_log_set(C.cast(cb, C.c_void_p), None)
    try:
        with pytest.raises(api.WhisperError):
            WhisperContext.new_with_params("/nonexistent/model.bin", WhisperContextParameters(), lib=lib)
        with pytest.raises(api.WhisperError):
            WhisperContext.new_from_buffer_with_params(b"not a ggml file at all", WhisperContextParameters(), lib=lib)
        if not torch.cuda.is_available():
            with pytest.raises(api.WhisperError):
                WhisperContext.new_from_buffer_with_params(synth.model_bytes("micro"), WhisperContextParameters(), lib=lib)
            assert any(b"no CUDA device" in m or b"CPU path" in m for _, m in logs), logs[-3:]
    finally:
        api.install_quiet_logging(lib)


def test_use_gpu_false_is_refused_with_a_log_line():
    """there is no CPU path: a context asked for with use_gpu = false is not silently run on the GPU"""
    lib = api.load_library()
    logs = []
    cb = api.LOG_CB(lambda lvl, msg, ud: logs.append((lvl, msg)))
    lib.whisper_log_set(C.cast(cb, C.c_void_p), None)
    try:
        with pytest.raises(api.WhisperError):
            WhisperContext.new_from_buffer_with_params(synth.model_bytes("micro", with_tensors=False), WhisperContextParameters(use_gpu=False), lib=lib)
        assert any(b"use_gpu = false" in m for _, m in logs), logs[-3:]
    finally:
        api.install_quiet_logging(lib)
