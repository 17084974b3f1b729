"""Host-side mirror of the reference's safe API for the transcription path, over the
whisper.h C ABI (the drop-in boundary).

The reference's host layer is Rust (`src/whisper_ctx_wrapper.rs`, `src/whisper_state.rs`,
`src/whisper_params.rs`); no Rust toolchain exists in this image, so the same surface —
same names, argument meaning and error behaviour — is restated here in Python over ctypes.
Every class takes the shared library to bind, so one binding drives both the product
library (`libwhisper_b200.so`) and, *in tests only*, the oracle build of the reference
(`oracle/_ref/libwhisper_ref.so`): the parity tests therefore read like the reference's own
examples (`examples/basic_use.rs:319-345`).

No arithmetic happens in this file.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from enum import IntEnum
from pathlib import Path

import numpy as np

_PKG = Path(__file__).resolve().parent
PRODUCT_LIB = _PKG / "lib" / "libwhisper_b200.so"


class WhisperError(Exception):
    """Mirrors `WhisperError` (src/error.rs:9-51): `code` carries the C return value."""

    def __init__(self, kind: str, code: int = 0):
        super().__init__(f"{kind} ({code})" if code else kind)
        self.kind, self.code = kind, code


# ----------------------------------------------------------------------------------------
# ABI structs (whisper.h:116-151, 485-588; sizes asserted against SURVEY.md §8b)
# ----------------------------------------------------------------------------------------
class whisper_ahead(C.Structure):
    _fields_ = [("n_text_layer", C.c_int), ("n_head", C.c_int)]


class whisper_aheads(C.Structure):
    _fields_ = [("n_heads", C.c_size_t), ("heads", C.POINTER(whisper_ahead))]


class whisper_context_params(C.Structure):
    _fields_ = [("use_gpu", C.c_bool), ("flash_attn", C.c_bool), ("gpu_device", C.c_int),
                ("dtw_token_timestamps", C.c_bool), ("dtw_aheads_preset", C.c_int), ("dtw_n_top", C.c_int),
                ("dtw_aheads", whisper_aheads), ("dtw_mem_size", C.c_size_t)]


class whisper_token_data(C.Structure):
    _fields_ = [("id", C.c_int32), ("tid", C.c_int32), ("p", C.c_float), ("plog", C.c_float),
                ("pt", C.c_float), ("ptsum", C.c_float), ("t0", C.c_int64), ("t1", C.c_int64),
                ("t_dtw", C.c_int64), ("vlen", C.c_float)]


class whisper_vad_params(C.Structure):
    _fields_ = [("threshold", C.c_float), ("min_speech_duration_ms", C.c_int), ("min_silence_duration_ms", C.c_int),
                ("max_speech_duration_s", C.c_float), ("speech_pad_ms", C.c_int), ("samples_overlap", C.c_float)]


class _greedy(C.Structure):
    _fields_ = [("best_of", C.c_int)]


class _beam(C.Structure):
    _fields_ = [("beam_size", C.c_int), ("patience", C.c_float)]


NEW_SEGMENT_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p)
PROGRESS_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p)
ENCODER_BEGIN_CB = C.CFUNCTYPE(C.c_bool, C.c_void_p, C.c_void_p, C.c_void_p)
ABORT_CB = C.CFUNCTYPE(C.c_bool, C.c_void_p)
LOGITS_FILTER_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.POINTER(whisper_token_data), C.c_int,
                               C.POINTER(C.c_float), C.c_void_p)
LOG_CB = C.CFUNCTYPE(None, C.c_int, C.c_char_p, C.c_void_p)


class whisper_full_params(C.Structure):
    _fields_ = [
        ("strategy", C.c_int), ("n_threads", C.c_int), ("n_max_text_ctx", C.c_int), ("offset_ms", C.c_int),
        ("duration_ms", C.c_int),
        ("translate", C.c_bool), ("no_context", C.c_bool), ("no_timestamps", C.c_bool), ("single_segment", C.c_bool),
        ("print_special", C.c_bool), ("print_progress", C.c_bool), ("print_realtime", C.c_bool),
        ("print_timestamps", C.c_bool),
        ("token_timestamps", C.c_bool), ("thold_pt", C.c_float), ("thold_ptsum", C.c_float), ("max_len", C.c_int),
        ("split_on_word", C.c_bool), ("max_tokens", C.c_int),
        ("debug_mode", C.c_bool), ("audio_ctx", C.c_int), ("tdrz_enable", C.c_bool),
        ("suppress_regex", C.c_char_p), ("initial_prompt", C.c_char_p),
        ("prompt_tokens", C.POINTER(C.c_int32)), ("prompt_n_tokens", C.c_int),
        ("language", C.c_char_p), ("detect_language", C.c_bool),
        ("suppress_blank", C.c_bool), ("suppress_nst", C.c_bool),
        ("temperature", C.c_float), ("max_initial_ts", C.c_float), ("length_penalty", C.c_float),
        ("temperature_inc", C.c_float), ("entropy_thold", C.c_float), ("logprob_thold", C.c_float),
        ("no_speech_thold", C.c_float),
        ("greedy", _greedy), ("beam_search", _beam),
        ("new_segment_callback", NEW_SEGMENT_CB), ("new_segment_callback_user_data", C.c_void_p),
        ("progress_callback", PROGRESS_CB), ("progress_callback_user_data", C.c_void_p),
        ("encoder_begin_callback", ENCODER_BEGIN_CB), ("encoder_begin_callback_user_data", C.c_void_p),
        ("abort_callback", ABORT_CB), ("abort_callback_user_data", C.c_void_p),
        ("logits_filter_callback", LOGITS_FILTER_CB), ("logits_filter_callback_user_data", C.c_void_p),
        ("grammar_rules", C.c_void_p), ("n_grammar_rules", C.c_size_t), ("i_start_rule", C.c_size_t),
        ("grammar_penalty", C.c_float),
        ("vad", C.c_bool), ("vad_model_path", C.c_char_p), ("vad_params", whisper_vad_params),
    ]


assert C.sizeof(whisper_full_params) == 296 and C.sizeof(whisper_context_params) == 48
assert C.sizeof(whisper_token_data) == 56
assert whisper_full_params.language.offset == 96 and whisper_full_params.greedy.offset == 136
assert whisper_full_params.new_segment_callback.offset == 152 and whisper_full_params.vad.offset == 260


class SamplingStrategy(IntEnum):
    """`whisper_sampling_strategy` (whisper.h:453-456) / `SamplingStrategy` (src/whisper_params.rs)."""
    GREEDY = 0
    BEAM_SEARCH = 1


class DtwModelPreset(IntEnum):
    """`whisper_alignment_heads_preset` (whisper.h:88-104)."""
    NONE = 0
    N_TOP_MOST = 1
    CUSTOM = 2
    TINY_EN = 3
    TINY = 4
    BASE_EN = 5
    BASE = 6
    SMALL_EN = 7
    SMALL = 8
    MEDIUM_EN = 9
    MEDIUM = 10
    LARGE_V1 = 11
    LARGE_V2 = 12
    LARGE_V3 = 13
    LARGE_V3_TURBO = 14


_V, _I, _P = C.c_void_p, C.c_int, C.c_char_p
_SIGS = {
    # name: (restype, argtypes)
    "whisper_context_default_params": (whisper_context_params, []),
    "whisper_full_default_params": (whisper_full_params, [_I]),
    "whisper_init_from_file_with_params_no_state": (_V, [_P, whisper_context_params]),
    "whisper_init_from_buffer_with_params_no_state": (_V, [_V, C.c_size_t, whisper_context_params]),
    "whisper_init_state": (_V, [_V]),
    "whisper_free": (None, [_V]),
    "whisper_free_state": (None, [_V]),
    "whisper_full_with_state": (_I, [_V, _V, whisper_full_params, C.POINTER(C.c_float), _I]),
    "whisper_pcm_to_mel_with_state": (_I, [_V, _V, C.POINTER(C.c_float), _I, _I]),
    "whisper_set_mel_with_state": (_I, [_V, _V, C.POINTER(C.c_float), _I, _I]),
    "whisper_encode_with_state": (_I, [_V, _V, _I, _I]),
    "whisper_decode_with_state": (_I, [_V, _V, C.POINTER(C.c_int32), _I, _I, _I]),
    "whisper_lang_auto_detect_with_state": (_I, [_V, _V, _I, _I, C.POINTER(C.c_float)]),
    "whisper_get_logits_from_state": (C.POINTER(C.c_float), [_V]),
    "whisper_full_n_segments_from_state": (_I, [_V]),
    "whisper_full_lang_id_from_state": (_I, [_V]),
    "whisper_full_get_segment_t0_from_state": (C.c_int64, [_V, _I]),
    "whisper_full_get_segment_t1_from_state": (C.c_int64, [_V, _I]),
    "whisper_full_get_segment_text_from_state": (_P, [_V, _I]),
    "whisper_full_get_segment_speaker_turn_next_from_state": (C.c_bool, [_V, _I]),
    "whisper_full_get_segment_no_speech_prob_from_state": (C.c_float, [_V, _I]),
    "whisper_full_n_tokens_from_state": (_I, [_V, _I]),
    "whisper_full_get_token_text_from_state": (_P, [_V, _V, _I, _I]),
    "whisper_full_get_token_id_from_state": (C.c_int32, [_V, _I, _I]),
    "whisper_full_get_token_data_from_state": (whisper_token_data, [_V, _I, _I]),
    "whisper_full_get_token_p_from_state": (C.c_float, [_V, _I, _I]),
    "whisper_n_len_from_state": (_I, [_V]),
    "whisper_n_vocab": (_I, [_V]), "whisper_n_text_ctx": (_I, [_V]), "whisper_n_audio_ctx": (_I, [_V]),
    "whisper_is_multilingual": (_I, [_V]),
    "whisper_model_n_vocab": (_I, [_V]), "whisper_model_n_audio_ctx": (_I, [_V]),
    "whisper_model_n_audio_state": (_I, [_V]), "whisper_model_n_audio_head": (_I, [_V]),
    "whisper_model_n_audio_layer": (_I, [_V]), "whisper_model_n_text_ctx": (_I, [_V]),
    "whisper_model_n_text_state": (_I, [_V]), "whisper_model_n_text_head": (_I, [_V]),
    "whisper_model_n_text_layer": (_I, [_V]), "whisper_model_n_mels": (_I, [_V]),
    "whisper_model_ftype": (_I, [_V]), "whisper_model_type": (_I, [_V]),
    "whisper_model_type_readable": (_P, [_V]),
    "whisper_tokenize": (_I, [_V, _P, C.POINTER(C.c_int32), _I]),
    "whisper_token_to_str": (_P, [_V, C.c_int32]),
    "whisper_token_eot": (C.c_int32, [_V]), "whisper_token_sot": (C.c_int32, [_V]),
    "whisper_token_solm": (C.c_int32, [_V]), "whisper_token_prev": (C.c_int32, [_V]),
    "whisper_token_nosp": (C.c_int32, [_V]), "whisper_token_not": (C.c_int32, [_V]),
    "whisper_token_beg": (C.c_int32, [_V]), "whisper_token_lang": (C.c_int32, [_V, _I]),
    "whisper_token_translate": (C.c_int32, [_V]), "whisper_token_transcribe": (C.c_int32, [_V]),
    "whisper_lang_id": (_I, [_P]), "whisper_lang_max_id": (_I, []),
    "whisper_lang_str": (_P, [_I]), "whisper_lang_str_full": (_P, [_I]),
    "whisper_log_set": (None, [_V, _V]),
    "whisper_print_system_info": (_P, []),
}


def load_library(path: os.PathLike | str | None = None) -> C.CDLL:
    """dlopen a library exporting the whisper.h ABI and declare the signatures used here.
    Default is the product library; it must exist (there is no fallback)."""
    p = Path(path) if path is not None else PRODUCT_LIB
    if not p.exists():
        raise WhisperError(f"shared library not found: {p} (build it with __graft_entry__.build())")
    lib = C.CDLL(str(p), mode=C.RTLD_LOCAL)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    lib._path = str(p)
    return lib


@dataclass
class DtwParameters:
    """`DtwParameters` (src/whisper_ctx.rs:472-657)."""
    preset: DtwModelPreset = DtwModelPreset.NONE
    n_top: int = -1
    custom_heads: tuple = ()
    dtw_mem_size: int = 1024 * 1024 * 128


@dataclass
class WhisperContextParameters:
    """`WhisperContextParameters` (src/whisper_ctx.rs:472-520): defaults as `whisper_context_default_params`."""
    use_gpu: bool = True
    flash_attn: bool = False
    gpu_device: int = 0
    dtw_parameters: DtwParameters | None = None

    def to_c(self, lib):
        cp = lib.whisper_context_default_params()
        cp.use_gpu, cp.flash_attn, cp.gpu_device = self.use_gpu, self.flash_attn, self.gpu_device
        keep = None
        d = self.dtw_parameters
        if d is not None and d.preset != DtwModelPreset.NONE:
            cp.dtw_token_timestamps = True
            cp.dtw_aheads_preset = int(d.preset)
            cp.dtw_n_top = d.n_top
            cp.dtw_mem_size = d.dtw_mem_size
            if d.preset == DtwModelPreset.CUSTOM:
                keep = (whisper_ahead * len(d.custom_heads))(*[whisper_ahead(l, h) for l, h in d.custom_heads])
                cp.dtw_aheads = whisper_aheads(len(d.custom_heads), keep)
        return cp, keep


class FullParams:
    """`FullParams` (src/whisper_params.rs:48-802): a `whisper_full_params` built from
    `whisper_full_default_params(strategy)` plus setters of the same names."""

    def __init__(self, lib, strategy: SamplingStrategy = SamplingStrategy.GREEDY, best_of: int | None = None,
                 beam_size: int | None = None, patience: float = -1.0):
        self._lib = lib
        self.fp = lib.whisper_full_default_params(int(strategy))
        if strategy == SamplingStrategy.GREEDY and best_of is not None:
            self.fp.greedy.best_of = best_of
        if strategy == SamplingStrategy.BEAM_SEARCH:
            if beam_size is not None:
                self.fp.beam_search.beam_size = beam_size
            self.fp.beam_search.patience = patience
        self._keep = {}
        # the reference examples switch the printing off (examples/basic_use.rs:326-333)
        self.fp.print_progress = False
        self.fp.print_realtime = False
        self.fp.print_special = False
        self.fp.print_timestamps = False

    def __getattr__(self, name):
        if name.startswith("set_"):
            field = name[4:]

            def setter(v):
                if field in ("language", "initial_prompt", "suppress_regex"):
                    b = None if v is None else (v if isinstance(v, bytes) else str(v).encode())
                    self._keep[field] = b
                    setattr(self.fp, field, b)
                elif field == "tokens":
                    arr = (C.c_int32 * len(v))(*v)
                    self._keep[field] = arr
                    self.fp.prompt_tokens = C.cast(arr, C.POINTER(C.c_int32))
                    self.fp.prompt_n_tokens = len(v)
                elif field.endswith("_callback"):
                    typ = dict(new_segment_callback=NEW_SEGMENT_CB, progress_callback=PROGRESS_CB,
                               encoder_begin_callback=ENCODER_BEGIN_CB, abort_callback=ABORT_CB,
                               logits_filter_callback=LOGITS_FILTER_CB)[field]
                    cb = typ(v) if v is not None else typ()
                    self._keep[field] = cb
                    setattr(self.fp, field, cb)
                else:
                    setattr(self.fp, field, v)
                return self
            return setter
        raise AttributeError(name)


@dataclass
class Segment:
    t0: int
    t1: int
    text: bytes
    tokens: list   # list[whisper_token_data-like dict]
    no_speech_prob: float = 0.0


class WhisperContext:
    """`WhisperContext` (src/whisper_ctx_wrapper.rs:35-454)."""

    def __init__(self, lib, ptr, keep=None):
        self.lib, self.ptr, self._keep = lib, ptr, keep

    @classmethod
    def new_with_params(cls, path, params: WhisperContextParameters | None = None, lib=None):
        lib = lib or load_library()
        cp, keep = (params or WhisperContextParameters()).to_c(lib)
        ptr = lib.whisper_init_from_file_with_params_no_state(os.fspath(path).encode(), cp)
        if not ptr:
            raise WhisperError("InitError")
        return cls(lib, ptr, keep)

    @classmethod
    def new_from_buffer_with_params(cls, buf: bytes, params: WhisperContextParameters | None = None, lib=None):
        lib = lib or load_library()
        cp, keep = (params or WhisperContextParameters()).to_c(lib)
        cbuf = (C.c_char * len(buf)).from_buffer_copy(buf)
        ptr = lib.whisper_init_from_buffer_with_params_no_state(C.cast(cbuf, C.c_void_p), len(buf), cp)
        if not ptr:
            raise WhisperError("InitError")
        return cls(lib, ptr, keep)

    def create_state(self) -> "WhisperState":
        st = self.lib.whisper_init_state(self.ptr)
        if not st:
            raise WhisperError("InitError")
        return WhisperState(self, st)

    # model / tokenizer getters (src/whisper_ctx.rs:83-457)
    def n_vocab(self): return self.lib.whisper_n_vocab(self.ptr)
    def n_text_ctx(self): return self.lib.whisper_n_text_ctx(self.ptr)
    def n_audio_ctx(self): return self.lib.whisper_n_audio_ctx(self.ptr)
    def is_multilingual(self): return self.lib.whisper_is_multilingual(self.ptr) != 0
    def model_n_mels(self): return self.lib.whisper_model_n_mels(self.ptr)
    def model_type_readable(self): return self.lib.whisper_model_type_readable(self.ptr)
    def token_eot(self): return self.lib.whisper_token_eot(self.ptr)
    def token_sot(self): return self.lib.whisper_token_sot(self.ptr)
    def token_beg(self): return self.lib.whisper_token_beg(self.ptr)
    def token_not(self): return self.lib.whisper_token_not(self.ptr)
    def token_lang(self, i): return self.lib.whisper_token_lang(self.ptr, i)
    def token_transcribe(self): return self.lib.whisper_token_transcribe(self.ptr)
    def token_to_str(self, t): return self.lib.whisper_token_to_str(self.ptr, t)

    def tokenize(self, text: str, max_tokens: int) -> list:
        buf = (C.c_int32 * max_tokens)()
        n = self.lib.whisper_tokenize(self.ptr, text.encode(), buf, max_tokens)
        if n < 0:
            raise WhisperError("InvalidText", n)
        return list(buf[:n])

    def free(self):
        if self.ptr:
            self.lib.whisper_free(self.ptr)
            self.ptr = None


class WhisperState:
    """`WhisperState` (src/whisper_state.rs:14-607)."""

    def __init__(self, ctx: WhisperContext, ptr):
        self.ctx, self.lib, self.ptr = ctx, ctx.lib, ptr

    def free(self):
        if self.ptr:
            self.lib.whisper_free_state(self.ptr)
            self.ptr = None

    @staticmethod
    def _f32(a):
        a = np.ascontiguousarray(a, dtype=np.float32)
        return a, a.ctypes.data_as(C.POINTER(C.c_float))

    def pcm_to_mel(self, pcm, threads: int = 1):
        a, p = self._f32(pcm)
        if self.lib.whisper_pcm_to_mel_with_state(self.ctx.ptr, self.ptr, p, len(a), threads) != 0:
            raise WhisperError("UnableToCalculateSpectrogram")

    def set_mel(self, mel: np.ndarray):
        """mel: [n_mel, n_len] float32 (mel-major, as the engine stores it)."""
        a, p = self._f32(mel)
        r = self.lib.whisper_set_mel_with_state(self.ctx.ptr, self.ptr, p, a.shape[1], a.shape[0])
        if r != 0:
            raise WhisperError("InvalidMelBands", r)

    def encode(self, offset: int = 0, threads: int = 1):
        r = self.lib.whisper_encode_with_state(self.ctx.ptr, self.ptr, offset, threads)
        if r != 0:
            raise WhisperError("UnableToCalculateEvaluation", r)

    def decode(self, tokens, n_past: int, threads: int = 1):
        arr = (C.c_int32 * len(tokens))(*tokens)
        r = self.lib.whisper_decode_with_state(self.ctx.ptr, self.ptr, arr, len(tokens), n_past, threads)
        if r != 0:
            raise WhisperError("UnableToCalculateEvaluation", r)

    def get_logits(self) -> np.ndarray:
        p = self.lib.whisper_get_logits_from_state(self.ptr)
        if not p:
            raise WhisperError("NullPointer")
        return np.ctypeslib.as_array(p, shape=(self.ctx.n_vocab(),)).copy()

    def lang_detect(self, offset_ms: int = 0, threads: int = 1):
        n = self.lib.whisper_lang_max_id() + 1
        probs = (C.c_float * n)()
        r = self.lib.whisper_lang_auto_detect_with_state(self.ctx.ptr, self.ptr, offset_ms, threads, probs)
        if r < 0:
            raise WhisperError("GenericError", r)
        return r, np.array(probs[:], dtype=np.float32)

    def full(self, params: FullParams, pcm) -> int:
        """`WhisperState::full` (src/whisper_state.rs:289-321), incl. its return-code mapping."""
        a, p = self._f32(pcm)
        if len(a) == 0:
            raise WhisperError("NoSamples")
        r = self.lib.whisper_full_with_state(self.ctx.ptr, self.ptr, params.fp, p, len(a))
        if r == 0:
            return 0
        if r == 7:
            raise WhisperError("FailedToEncode", r)
        if r == 8:
            raise WhisperError("FailedToDecode", r)
        raise WhisperError("GenericError", r)

    def n_len(self): return self.lib.whisper_n_len_from_state(self.ptr)
    def full_n_segments(self): return self.lib.whisper_full_n_segments_from_state(self.ptr)
    def full_lang_id(self): return self.lib.whisper_full_lang_id_from_state(self.ptr)
    def full_get_segment_t0(self, i): return self.lib.whisper_full_get_segment_t0_from_state(self.ptr, i)
    def full_get_segment_t1(self, i): return self.lib.whisper_full_get_segment_t1_from_state(self.ptr, i)
    def full_get_segment_text(self, i): return self.lib.whisper_full_get_segment_text_from_state(self.ptr, i)
    def full_n_tokens(self, i): return self.lib.whisper_full_n_tokens_from_state(self.ptr, i)
    def full_get_token_id(self, i, j): return self.lib.whisper_full_get_token_id_from_state(self.ptr, i, j)
    def full_get_token_text(self, i, j): return self.lib.whisper_full_get_token_text_from_state(self.ctx.ptr, self.ptr, i, j)
    def full_get_token_prob(self, i, j): return self.lib.whisper_full_get_token_p_from_state(self.ptr, i, j)

    def full_get_token_data(self, i, j) -> dict:
        d = self.lib.whisper_full_get_token_data_from_state(self.ptr, i, j)
        return {k: getattr(d, k) for k, _ in whisper_token_data._fields_}

    def segments(self) -> list:
        out = []
        for i in range(self.full_n_segments()):
            toks = [self.full_get_token_data(i, j) for j in range(self.full_n_tokens(i))]
            out.append(Segment(self.full_get_segment_t0(i), self.full_get_segment_t1(i),
                               self.full_get_segment_text(i), toks,
                               self.lib.whisper_full_get_segment_no_speech_prob_from_state(self.ptr, i)))
        return out


def install_quiet_logging(lib):
    """Route the library's log callback to nothing (the reference installs `log`/`tracing`
    trampolines here: src/whisper_logging_hook.rs:17-73)."""
    cb = LOG_CB(lambda lvl, msg, ud: None)
    lib.whisper_log_set(C.cast(cb, C.c_void_p), None)
    lib._log_cb = cb
    return cb
